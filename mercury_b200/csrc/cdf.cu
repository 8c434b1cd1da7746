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
#include <cooperative_groups.h>

#include <atomic>
#include <math.h>

#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

#include "chain.cuh"
#include "common.cuh"
#include "internal.cuh"
#include "score_small.cuh"

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
// phase guesses (used only to choose which maps to precompute; a wrong guess costs time, not bits)
constexpr uint32_t kAnyPhase = 0x3ffu;    // all addends zero: the identity map in every phase
constexpr uint32_t kPhaseMask = 0x3ffu;
constexpr uint32_t kFlag = 0x80000000u;   // chunk: pack present; sub-chunk: sum it serially (phase change expected)
constexpr uint32_t kTie0 = 0x40000000u;   // the map for the guessed phase is not a plain increment (exact ties inside)
constexpr uint32_t kTie1 = 0x20000000u;   // ... nor is the map for the phase above
constexpr uint32_t kNoPack = 0x10000000u; // chunk expected to change phase but no pack slot was left for it
constexpr int kSub = 64, kSubsPerChunk = kChunk / kSub;
constexpr int kSubSlots = 128;            // flagged chunks that get a pack (compact layout)
constexpr int kRawPerPack = 6;
// Margins around the guesses.  The fp32 chain drifts away from the exact running total (the
// guesses come from the exact one): ~1e-5 relative at a million candidates, up to several 1e-3 at
// 4-8 million, where an addend is only a few ulps of the sum.  A flagged chunk costs a pack
// (~1.5k cycles of the walk), a flagged sub-chunk a serial pass (~450 cycles): generous margins
// only where they are needed.
__host__ __device__ inline double chunk_margin(int nchunks) { return nchunks <= 1024 ? 1.0 / 4096.0 : 1.0 / 1024.0; }
__host__ __device__ inline double sub_margin(int nchunks) { return nchunks <= 1024 ? 1.0 / 16384.0 : 1.0 / 8192.0; }
static_assert(kSubsPerChunk == 32, "one sub-chunk per lane of the walking warp");

// Guess the phase of the running sum over a span from its approximate value at both ends, with
// a margin for the drift of the fp32 chain: returns the low guess, and whether the span may
// change phase.
__device__ __forceinline__ bool guess_span(double a_lo, double a_hi, double bias, uint32_t* e_lo, double margin) {
    const uint32_t lo = phase_of(f2b((float)(a_lo * bias * (1.0 - margin))));
    const uint32_t hi = phase_of(f2b((float)(a_hi * bias * (1.0 + margin))));
    *e_lo = lo;
    return lo != hi;
}

// Exclusive scan (fp64) of one value per chunk over a 1024-thread CTA -> approx[0..nchunks], and
// a compact slot number for every chunk that may change phase -> slotid[c] (-1: none).
template <typename F>
__device__ __forceinline__ void scan_chunks(F value_of, int nchunks, double bias, int max_slots, double* __restrict__ approx,
                                            int* __restrict__ slotid, double* red, int* redi) {
    __shared__ double carry;
    __shared__ int carry_slots;
    if (threadIdx.x == 0) { carry = 0.0; carry_slots = 0; }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c0 = 0; c0 < nchunks; c0 += 1024) {
        const int c = c0 + threadIdx.x;
        const double v = c < nchunks ? value_of(c) : 0.0;
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
        const double a_lo = pre + inc - v, a_hi = pre + inc;
        uint32_t e;
        const bool flagged = c < nchunks && guess_span(a_lo, a_hi, bias, &e, chunk_margin(nchunks));
        const unsigned bal = __ballot_sync(0xffffffffu, flagged);
        if (lane == 0) redi[warp] = __popc(bal);
        __syncthreads();
        int before = carry_slots;
        for (int i = 0; i < warp; ++i) before += redi[i];
        before += __popc(bal & ((1u << lane) - 1u));
        if (c < nchunks) {
            approx[c] = a_lo;
            slotid[c] = flagged ? (before < max_slots ? before : -2) : -1;  // -2: flagged, but no slot left
        }
        __syncthreads();
        if (threadIdx.x == 1023) { carry = a_hi; carry_slots = before + (flagged ? 1 : 0); }
        __syncthreads();
    }
    if (threadIdx.x == 0) approx[nchunks] = carry;
}

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
            if (i >= pos && i < len) loc = compose(loc, addend_pair_fast(pb[k], E));
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
                const Pair d = addend_pair_fast(pb[k], E);
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
__device__ __forceinline__ void load_chunk(const WeightSrc& src, int64_t start, int len, uint32_t (&pb)[IPT], float shift,
                                           float total) {
    const int base = threadIdx.x * IPT;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
        const int i = base + k;
        pb[k] = i < len ? f2b(src.get(start + i, shift, total)) : 0u;
    }
}
template <int NT, int IPT>
__device__ __forceinline__ void load_chunk(const WeightSrc& src, int64_t start, int len, uint32_t (&pb)[IPT]) {
    float shift, total;
    src.prepare(shift, total);
    load_chunk<NT, IPT>(src, start, len, pb, shift, total);
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
                                                              double* __restrict__ approx, int* __restrict__ slotid, double bias,
                                                              int max_slots) {
    __shared__ double red[32];
    __shared__ int redi[32];
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
    }
    // exclusive scan of the per-chunk totals of p ~ (sum_scores + len*shift) / total
    scan_chunks([&](int c) {
        const int len = (int)min((int64_t)kChunk, n - (int64_t)c * kChunk);
        return (sums[c] + (double)len * (double)shift) / (double)total;
    }, nchunks, bias, max_slots, approx, slotid, red, redi);
}

// 2b. the same scan for raw weights
__global__ void __launch_bounds__(1024) prefix_finalize_kernel(const double* __restrict__ sums, int nchunks,
                                                               double* __restrict__ approx, int* __restrict__ slotid, double bias,
                                                              int max_slots) {
    __shared__ double red[32];
    __shared__ int redi[32];
    scan_chunks([&](int c) { return sums[c]; }, nchunks, bias, max_slots, approx, slotid, red, redi);
}

__global__ void __launch_bounds__(256) normalize_kernel(const float* __restrict__ scores, int64_t n,
                                                        const float* __restrict__ stats, float* __restrict__ p) {
    const float shift = stats[MERCURY_STAT_SHIFT], total = stats[MERCURY_STAT_TOTAL];
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        p[i] = __fdiv_rn(__fadd_rn(scores[i], shift), total);
}

// ---------------------------------------------------------------------------------------------
// 3. summaries: one map per chunk (for the guessed phase; and the next one where the running
//    sum may change phase), plus, for those "flagged" chunks, one map per 64-candidate sub-chunk
struct __align__(32) Summary { uint32_t e; Pair q[2]; uint32_t pad[3]; };  // e: guess | flags; pad[0]: pack index

// what the spine needs to cross a flagged chunk without scanning: per sub-chunk the guessed phase
// (| flags), the maps for that phase and the next, the exclusive prefix sums of their increments,
// and the raw candidates of (up to kRawPerPack) sub-chunks in which the phase should change
struct __align__(128) Pack {
    uint32_t e[kSubsPerChunk];
    Pair q0[kSubsPerChunk], q1[kSubsPerChunk];
    uint32_t ps0[kSubsPerChunk + 1], ps1[kSubsPerChunk + 1];
    uint8_t rawsub[8];                // sub-chunks whose raw candidates are in raw[] (0xff: unused)
    uint32_t nev;                     // the walk through the chunk, precompiled (see EvKind)
    uint32_t pad_[3];
    uint32_t raw[kRawPerPack][kSub];  // 16-byte aligned: read with 128-bit loads
    uint4 ev[kSubsPerChunk + 1];
};
static_assert(sizeof(Pack) == 3072 && offsetof(Pack, raw) % 16 == 0 && offsetof(Pack, ev) % 16 == 0, "Pack layout");

// One step of a precompiled walk.  x = kind | guard phase << 8 | first sub-chunk << 20 (packs).
//   kEvAdd   : the span is a plain increment while the running sum is in the guard phase: y (z: high word at chunk level)
//   kEvMap   : a span with exact ties inside: 2-state map (z, w)
//   kEvSerial: a sub-chunk in which the phase may change: x also carries its guessed phase, z / w its plain increments for
//              that phase and the next (y bits 8 / 9: not plain, exact ties inside); it is summed serially in fp32 only if
//              neither increment keeps the sum inside its phase.  y low byte = raw slot in the pack (0xff: fetch it)
//   kEvPack  : (chunk level) a flagged chunk: y = shared-memory slot of its pack (0xffffffff: fetch it)
//   kEvSlow  : (chunk level) a chunk that needs the general code
enum EvKind : uint32_t { kEvAdd = 0, kEvMap = 1, kEvSerial = 2, kEvPack = 3, kEvSlow = 4, kEvEnd = 5, kEvNone = 6 };  // (kEvEnd / kEvNone: step_fused.cuh)

std::atomic<double> g_guess_bias{1.0};  // test hook (mercury_test_set_guess_bias): scales every guess

constexpr int kSumNT = 256, kSumIPT = kChunk / kSumNT;
static_assert(kSumIPT * 8 == kSub, "a sub-chunk is 8 consecutive threads");

constexpr int kMaxShards = 64;
// ---- the same two steps with the exchange inside the kernel: every rank owns one PeerBox in
// memory its peers have mapped (CUDA IPC over NVLink / NVSwitch).  A rank stores its value, tagged
// with the step's sequence number in the same 64-bit word, straight into slot [rank] of every
// peer's box and then polls its OWN box until all G slots carry this step's tag: one NVLink store
// latency instead of a collective launch.  A word is (payload32 << 32 | seq32): written by one
// 8-byte store, so a reader never sees a payload without its tag and no fence is needed.  Slots
// are double-buffered by the tag's parity; a rank cannot be more than one exchange ahead of a peer
// (it needs that peer's next word to get any further).  Waits are bounded (kPeerWaitNs) and
// report MERCURY_ST_PEER_TIMEOUT instead of hanging the device.
struct PeerBox {
    unsigned long long sum_hi[2][kMaxShards];
    unsigned long long sum_lo[2][kMaxShards];
    unsigned long long total[2][kMaxShards];
    unsigned int seq_sum, seq_total;      // local step counters (peers never touch them)
    unsigned int dead;                    // set by the first timed-out wait: later waits give up at once (mercury_peer_reset clears it)
    unsigned long long wait_ns;           // bound of one wait (mercury_peer_create: MERCURY_PEER_WAIT_MS or 4 s; mercury_peer_set_wait)
};
static_assert(sizeof(PeerBox) <= MERCURY_PEER_BOX_BYTES, "PeerBox must fit the size the header promises");
constexpr unsigned long long kPeerWaitNs = 4000000000ull;   // default bound
// A step that saw a timeout must not look like a result: its draws get NaN probabilities and index -1, which survive the
// SUM all-reduce that assembles a batch, and MERCURY_ST_PEER_TIMEOUT is set for the host to raise on.
__device__ __forceinline__ double peer_poison() { return __longlong_as_double(0x7ff8000000000000ll); }

__device__ __forceinline__ void peer_store(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_load(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// polls *p until its low half equals seq; false on timeout
__device__ __forceinline__ bool peer_wait(const unsigned long long* p, unsigned int seq, unsigned long long& word, PeerBox* box) {
    volatile unsigned int* dead = &box->dead;
    word = peer_load(p);
    if ((unsigned int)word == seq) return true;
    if (*dead) return false;
    const unsigned long long bound = box->wait_ns ? box->wait_ns : kPeerWaitNs;
    const unsigned long long t0 = global_ns();
    for (;;) {
        for (int i = 0; i < 64; ++i) {
            word = peer_load(p);
            if ((unsigned int)word == seq) return true;
        }
        if (global_ns() - t0 > bound) {
            *dead = 1u;
            return false;
        }
    }
}


// Statistics folded into the summaries (pools of up to kFusedMaxChunks chunks, one EMA update):
// every CTA reduces the partial sums of the scores itself -- same data, same order, so all CTAs
// agree to the bit on mean / EMA / shift / total -- which removes a single-CTA kernel from the
// dependency chain.  CTA 0 publishes the statistics block; the spine stores the new EMA state.
constexpr int kFusedMaxChunks = 8192;
struct FusedStats {
    const double* part;  // sums of `per` consecutive scores each (per divides kChunk)
    int npart, per;
    const float* ema_in; int ema_first; double ema_alpha, smooth_alpha;
    float* stats;
    // sharded pools: this launch covers one shard, the statistics are those of the WHOLE pool --
    // gsums[0..ngsums) are the per-shard sums of scores (all-gathered), n_global the pool size
    const double* gsums; int ngsums; int64_t n_global;
    // ... or exchanged right here through the peers' boxes (the whole sharded step as one chain of four kernels):
    // CTA 0 pushes this shard's sum, every CTA waits for all G sums in its own box; gsums_out (fp64 [nshards]) is
    // filled by CTA 0 for the caller
    PeerBox* const* peers; int nshards, rank; double* gsums_out;
    // ... where only CTA 0 waits for the peers: it hands the whole pool's sum to the other CTAs in two words of this GPU's
    // memory, (half of the double's bits << 32 | the exchange's tag) -- every CTA polling the peers' words itself is a storm of
    // system-scope loads on one cache line that grows with the number of shards
    unsigned long long* gsum;
};

// PEER: sharded pool with the sums exchanged inside this kernel (a separate instantiation, so that the one-GPU kernel
// carries none of it)
template <bool FUSED, bool PEER = false>
__global__ void __launch_bounds__(kSumNT) summarize_kernel(WeightSrc src, int64_t n, const double* __restrict__ approx,
                                                           const int* __restrict__ slotid, double bias,
                                                           Summary* __restrict__ out, Pack* __restrict__ subs,
                                                           int32_t* status, int check, FusedStats fs) {
    __shared__ Pair wt[2][kSumNT / 32];
    __shared__ float wsum[kSumNT / 32];
    __shared__ int wnz[kSumNT / 32];
    __shared__ double dred[3][kSumNT / 32];
    pdl_wait();
    pdl_launch_dependents();
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t pb[kSumIPT];
    double a_lo, a_hi;
    int slot;
    if constexpr (FUSED) {
        // the chunk's raw scores first: their latency overlaps the reduction of the partial sums below
        float raw[kSumIPT];
        {
            const int base = threadIdx.x * kSumIPT;
            const float* wp = src.w + start + base;
            if (base + kSumIPT <= len && (reinterpret_cast<uintptr_t>(wp) & 15) == 0) {
#pragma unroll
                for (int k = 0; k < kSumIPT; k += 4) {
                    const float4 v = *reinterpret_cast<const float4*>(wp + k);
                    raw[k] = v.x; raw[k + 1] = v.y; raw[k + 2] = v.z; raw[k + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int k = 0; k < kSumIPT; ++k) raw[k] = base + k < len ? wp[k] : 0.f;
            }
        }
        const float ema_prev = *fs.ema_in;  // (early: its latency hides behind the loads below)
        PeerBox* const my_box = PEER ? fs.peers[fs.rank] : nullptr;
        const unsigned int sum_tag = PEER ? my_box->seq_sum + 1u : 0u;  // (the spine advances seq_sum once the step's sums are in)
        const int parts_per_chunk = kChunk / fs.per, first_part = (int)blockIdx.x * parts_per_chunk;   // (start / per without the 64-bit division)
        double tot = 0.0, pre = 0.0, own = 0.0;
        constexpr int kInFlight = 16;
        for (int i0 = threadIdx.x; i0 < fs.npart; i0 += kInFlight * kSumNT) {  // 16 loads in flight per thread
            double v[kInFlight];
#pragma unroll
            for (int k = 0; k < kInFlight; ++k) { const int i = i0 + k * kSumNT; v[k] = i < fs.npart ? fs.part[i] : 0.0; }
#pragma unroll
            for (int k = 0; k < kInFlight; ++k) {
                const int i = i0 + k * kSumNT;
                tot += v[k];
                if (i < first_part) pre += v[k];
                else if (i < first_part + parts_per_chunk) own += v[k];
            }
        }
        tot = warp_sum(tot); pre = warp_sum(pre); own = warp_sum(own);
        if (lane == 0) { dred[0][warp] = tot; dred[1][warp] = pre; dred[2][warp] = own; }
        __syncthreads();
        tot = pre = own = 0.0;
#pragma unroll
        for (int i = 0; i < kSumNT / 32; ++i) { tot += dred[0][i]; pre += dred[1][i]; own += dred[2][i]; }
        // pytorch_collab.py:109-112 -- mean, EMA (util.py:207-213), shift = alpha*ema, total = sum(s + shift)
        double n_stat = (double)n;
        if constexpr (PEER) {  // sharded pool, sums exchanged here by CTA 0; the other CTAs take the result from fs.gsum
            __shared__ double gs[kMaxShards];
            const int par = sum_tag & 1u, g = threadIdx.x;
            if (blockIdx.x == 0) {
                if (g < fs.nshards) {
                    const unsigned long long bits = (unsigned long long)__double_as_longlong(tot);
                    peer_store(&fs.peers[g]->sum_hi[par][fs.rank], (bits >> 32 << 32) | sum_tag);
                    peer_store(&fs.peers[g]->sum_lo[par][fs.rank], (bits << 32) | sum_tag);
                    unsigned long long hi, lo;
                    const bool ok = peer_wait(&my_box->sum_hi[par][g], sum_tag, hi, my_box) &
                                    peer_wait(&my_box->sum_lo[par][g], sum_tag, lo, my_box);
                    if (!ok) atomicOr(status, MERCURY_ST_PEER_TIMEOUT);
                    gs[g] = __longlong_as_double((long long)((hi >> 32 << 32) | (lo >> 32)));
                    fs.gsums_out[g] = gs[g];
                }
                __syncthreads();
                tot = 0.0;
                for (int i = 0; i < fs.nshards; ++i) tot += gs[i];   // (same order on every rank)
                if (threadIdx.x == 0) {
                    const unsigned long long bits = (unsigned long long)__double_as_longlong(tot);
                    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(fs.gsum), "l"((bits >> 32 << 32) | sum_tag) : "memory");
                    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(fs.gsum + 1), "l"((bits << 32) | sum_tag) : "memory");
                }
            } else {
                if (threadIdx.x == 0) {   // (bounded like a peer wait, twice as long: CTA 0 may itself be waiting for a peer)
                    unsigned long long hi = 0ull, lo = 0ull;
                    const unsigned long long bound = 2ull * (my_box->wait_ns ? my_box->wait_ns : kPeerWaitNs), t0 = global_ns();
                    for (unsigned int it = 0;; ++it) {
                        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(hi) : "l"(fs.gsum) : "memory");
                        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(lo) : "l"(fs.gsum + 1) : "memory");
                        if ((unsigned int)hi == sum_tag && (unsigned int)lo == sum_tag) break;
                        if ((it & 63u) == 63u && global_ns() - t0 > bound) { atomicOr(status, MERCURY_ST_PEER_TIMEOUT); break; }
                    }
                    gs[0] = __longlong_as_double((long long)((hi >> 32 << 32) | (lo >> 32)));
                }
                __syncthreads();
                tot = gs[0];
            }
            n_stat = (double)fs.n_global;
        } else if (fs.gsums) {  // whole-pool statistics from the shard sums (same order on every rank)
            tot = 0.0;
            for (int i = 0; i < fs.ngsums; ++i) tot += fs.gsums[i];
            n_stat = (double)fs.n_global;
        }
        const float mean = (float)(tot / n_stat);
        const float a = (float)fs.ema_alpha, one_minus_a = (float)(1.0 - fs.ema_alpha);
        const float ema = fs.ema_first ? mean : __fadd_rn(__fmul_rn(a, ema_prev), __fmul_rn(one_minus_a, mean));
        const float shift = __fmul_rn((float)fs.smooth_alpha, ema);
        const float total = (float)(tot + n_stat * (double)shift);
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            fs.stats[MERCURY_STAT_MEAN] = mean; fs.stats[MERCURY_STAT_EMA] = ema;
            fs.stats[MERCURY_STAT_SHIFT] = shift; fs.stats[MERCURY_STAT_TOTAL] = total;
            fs.stats[4] = fs.stats[5] = fs.stats[6] = fs.stats[7] = 0.f;
        }
        const double inv_total = 1.0 / (double)total;   // (one division: the approximate running totals only feed the guesses)
        a_lo = (pre + (double)start * (double)shift) * inv_total;
        a_hi = a_lo + (own + (double)len * (double)shift) * inv_total;
        {
            const int base = threadIdx.x * kSumIPT;
#pragma unroll
            for (int k = 0; k < kSumIPT; ++k) pb[k] = base + k < len ? f2b(__fdiv_rn(__fadd_rn(raw[k], shift), total)) : 0u;
        }
        slot = -2;  // decided below: flagged chunks index `subs` directly
    } else {
        a_lo = approx[blockIdx.x]; a_hi = approx[blockIdx.x + 1];
        slot = slotid[blockIdx.x];
        load_chunk<kSumNT, kSumIPT>(src, start, len, pb);
    }
    // Guess the phase from the approximate running total at both ends of the chunk.  If both ends
    // agree the chunk is expected to stay in one phase; otherwise it is flagged and (if a slot was
    // assigned to it) gets a pack of sub-chunk summaries for the spine.
    uint32_t e;
    const bool flagged = guess_span(a_lo, a_hi, bias, &e, chunk_margin((int)gridDim.x));  // block-uniform
    if (FUSED) slot = flagged ? (int)blockIdx.x : -1;
    Pair a0 = identity_pair();
    int bad = 0;
    bool nz = false;
    float tsum = 0.f;
#pragma unroll
    for (int k = 0; k < kSumIPT; ++k) {
        const float x = b2f(pb[k]);
        if (check) {
            if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
            else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
        }
        nz |= pb[k] != 0u;
        tsum += x;
        a0 = compose(a0, addend_pair_fast(pb[k], e));
    }
    if (bad) atomicOr(status, bad);
    const unsigned nzmask = __ballot_sync(0xffffffffu, nz);
    // ordered reduction: lane 0 ends with the composition of the whole warp.  Without exact ties
    // every map is a plain increment and the composition is one integer warp reduction.
    if (__ballot_sync(0xffffffffu, a0.d0 != a0.d1) == 0u) {
        a0.d0 = a0.d1 = sat(__reduce_add_sync(0xffffffffu, a0.d0));
    } else {
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            Pair o0;
            o0.d0 = __shfl_down_sync(0xffffffffu, a0.d0, off); o0.d1 = __shfl_down_sync(0xffffffffu, a0.d1, off);
            if (lane + off < 32) a0 = compose(a0, o0);
        }
    }
    float inc = tsum;  // inclusive scan of the thread sums (approximate, only feeds sub-chunk guesses)
    if (slot >= 0) {
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const float o = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += o;
        }
    }
    if (lane == 0) { wt[0][warp] = a0; wnz[warp] = nzmask != 0u; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        Pair t0 = identity_pair();
        int any = 0;
#pragma unroll
        for (int i = 0; i < kSumNT / 32; ++i) { t0 = compose(t0, wt[0][i]); any |= wnz[i]; }
        Summary s;
        s.e = (any ? e : kAnyPhase) | (slot >= 0 ? kFlag : 0u) | (flagged && slot < 0 ? kNoPack : 0u) | (t0.d0 != t0.d1 ? kTie0 : 0u);
        s.q[0] = t0;
        s.q[1] = Pair{kSatCap, kSatCap};  // (unused: flagged chunks are crossed through their packs)
        s.pad[0] = (uint32_t)slot; s.pad[1] = s.pad[2] = 0;
        out[blockIdx.x] = s;
    }
    if (slot < 0) return;
    // ---- per sub-chunk (8 consecutive threads = 64 candidates)
    float pre = 0.f;
    for (int i = 0; i < warp; ++i) pre += wsum[i];
    const float t_lo = pre + inc - tsum, t_hi = pre + inc;          // running total at this thread's ends
    const int lead = lane & ~7;
    const float s_lo = __shfl_sync(0xffffffffu, t_lo, lead), s_hi = __shfl_sync(0xffffffffu, t_hi, lead + 7);
    uint32_t es;
    // (a sub-chunk of zeros changes nothing, whatever the phase: never worth a serial pass)
    const bool sub_flagged = guess_span(a_lo + (double)s_lo, a_lo + (double)s_hi, bias, &es, sub_margin((int)gridDim.x)) &&
                             ((nzmask >> lead) & 0xffu) != 0u;
    Pair m0 = identity_pair(), m1 = identity_pair();
#pragma unroll
    for (int k = 0; k < kSumIPT; ++k) {
        m0 = compose(m0, addend_pair_fast(pb[k], es));
        m1 = compose(m1, addend_pair_fast(pb[k], es + 1u));
    }
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) {
        Pair o0, o1;
        o0.d0 = __shfl_down_sync(0xffffffffu, m0.d0, off, 8); o0.d1 = __shfl_down_sync(0xffffffffu, m0.d1, off, 8);
        o1.d0 = __shfl_down_sync(0xffffffffu, m1.d0, off, 8); o1.d1 = __shfl_down_sync(0xffffffffu, m1.d1, off, 8);
        if ((lane & 7) + off < 8) { m0 = compose(m0, o0); m1 = compose(m1, o1); }
    }
    __shared__ uint32_t sub_inc[2][kSubsPerChunk];
    __shared__ uint32_t sub_flag_mask;
    Pack* pk = subs + slot;
    const int my_sub = threadIdx.x >> 3;
    if ((lane & 7) == 0) {
        const bool sub_nz = ((nzmask >> lead) & 0xffu) != 0u;
        pk->e[my_sub] = (sub_nz ? es : kAnyPhase) | (sub_flagged ? kFlag : 0u) | (m0.d0 != m0.d1 ? kTie0 : 0u) |
                        (m1.d0 != m1.d1 ? kTie1 : 0u);
        pk->q0[my_sub] = m0;
        pk->q1[my_sub] = m1;
        sub_inc[0][my_sub] = m0.d0;
        sub_inc[1][my_sub] = m1.d0;
    }
    const unsigned lead_flags = __ballot_sync(0xffffffffu, (lane & 7) == 0 && sub_flagged);  // bits 0,8,16,24
    if (threadIdx.x == 0) sub_flag_mask = 0u;
    __syncthreads();
    if (lane == 0) {
        const uint32_t four = ((lead_flags >> 0) & 1u) | ((lead_flags >> 7) & 2u) | ((lead_flags >> 14) & 4u) | ((lead_flags >> 21) & 8u);
        atomicOr(&sub_flag_mask, four << (4 * warp));
    }
    __syncthreads();
    const uint32_t fmask = sub_flag_mask;
    if (warp == 0) {  // exclusive prefix sums of the increments over the 32 sub-chunks (32 * 2^25 fits)
        uint32_t v0 = sub_inc[0][lane], v1 = sub_inc[1][lane];
        const uint32_t i0 = v0, i1 = v1;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o0 = __shfl_up_sync(0xffffffffu, v0, off), o1 = __shfl_up_sync(0xffffffffu, v1, off);
            if (lane >= off) { v0 += o0; v1 += o1; }
        }
        pk->ps0[lane] = v0 - i0;
        pk->ps1[lane] = v1 - i1;
        if (lane == 31) { pk->ps0[32] = v0; pk->ps1[32] = v1; }
        // which sub-chunks ship their raw candidates: the first kRawPerPack flagged ones
        const int rank = __popc(fmask & ((1u << lane) - 1u));
        if (((fmask >> lane) & 1u) && rank < kRawPerPack) pk->rawsub[rank] = (uint8_t)lane;
        if (lane < kRawPerPack && lane >= __popc(fmask)) pk->rawsub[lane] = 0xff;
        // precompiled walk: maximal runs of plain sub-chunks -> one kEvAdd each; flagged -> kEvSerial; ties -> kEvMap
        const uint32_t ek = pk->e[lane];  // written above by this chunk's leaders (visible after the barrier)
        const uint32_t gk = ek & kPhaseMask;
        const unsigned any_m = __ballot_sync(0xffffffffu, gk == kAnyPhase);
        const unsigned tie_m = __ballot_sync(0xffffffffu, (ek & kTie0) != 0u) & ~fmask;
        const uint32_t gprev = __shfl_up_sync(0xffffffffu, gk, 1);
        const unsigned chg_m = __ballot_sync(0xffffffffu, lane > 0 && gk != kAnyPhase && gprev != kAnyPhase && gk != gprev);
        const unsigned stop_m = fmask | tie_m;
        const bool is_stop = (stop_m >> lane) & 1u;
        const bool run_start = !is_stop && (lane == 0 || ((stop_m >> (lane - 1)) & 1u) || ((chg_m >> lane) & 1u));
        const unsigned ev_m = __ballot_sync(0xffffffffu, is_stop || run_start);
        const unsigned brk = (stop_m | chg_m) & ~((2u << lane) - 1u);       // where a run starting here ends
        const int s_end = brk ? __ffs(brk) - 1 : kSubsPerChunk;
        const uint32_t ps_end = __shfl_sync(0xffffffffu, v0 - i0, s_end & 31);
        const uint32_t inc_run = (s_end == kSubsPerChunk ? __shfl_sync(0xffffffffu, v0, 31) : ps_end) - (v0 - i0);
        const unsigned span = (s_end == kSubsPerChunk ? 0xffffffffu : ((1u << s_end) - 1u)) & ~((1u << lane) - 1u);
        const unsigned live = span & ~any_m;                                 // first sub-chunk of the run with addends
        const uint32_t g_run = __shfl_sync(0xffffffffu, gk, live ? __ffs(live) - 1 : 0);
        if (is_stop || run_start) {
            uint4 ev;
            if (run_start) ev = make_uint4(kEvAdd | ((live ? g_run : kAnyPhase) << 8) | ((uint32_t)lane << 20), inc_run, 0u, 0u);
            else if ((fmask >> lane) & 1u) {
                // a flagged sub-chunk only MIGHT hold the phase change: carry its plain increments for the guessed phase
                // and the next, so that the walk sums it serially only when the change really falls inside
                const Pair f0 = pk->q0[lane], f1 = pk->q1[lane];
                ev = make_uint4(kEvSerial | (gk << 8) | ((uint32_t)lane << 20),
                                (rank < kRawPerPack ? (uint32_t)rank : 0xffu) | (f0.d0 != f0.d1 ? 0x100u : 0u) | (f1.d0 != f1.d1 ? 0x200u : 0u),
                                f0.d0, f1.d0);
            }
            else ev = make_uint4(kEvMap | (gk << 8) | ((uint32_t)lane << 20), 0u, pk->q0[lane].d0, pk->q0[lane].d1);
            pk->ev[__popc(ev_m & ((1u << lane) - 1u))] = ev;
        }
        if (lane == 0) pk->nev = __popc(ev_m);
    }
    const int my_rank = __popc(fmask & ((1u << my_sub) - 1u));
    if (((fmask >> my_sub) & 1u) && my_rank < kRawPerPack) {
#pragma unroll
        for (int k = 0; k < kSumIPT; ++k) pk->raw[my_rank][(threadIdx.x & 7) * kSumIPT + k] = pb[k];
    }
}

// ---------------------------------------------------------------------------------------------
// 4. the spine: exact running sum at every chunk boundary.
//
// The walk is serial by nature (about one step per change of binade, ~20 for a million
// candidates), so everything it needs is laid out beforehand, in parallel, such that a step costs
// a few dependent instructions instead of a warp scan:
//   * chunk level: while the running sum stays in the phase that was guessed, every chunk map is a
//     plain increment, so a whole run of chunks is crossed with ONE subtraction of prefix sums
//     (computed by the whole CTA while it loads the summaries into shared memory);
//   * a flagged chunk (phase change expected) carries a pack: per 64-candidate sub-chunk the maps
//     for the guessed phase and the next, their prefix sums, and the raw candidates of the
//     sub-chunks where the change should fall.  Runs of sub-chunks are again one subtraction; the
//     sub-chunk that changes phase is summed serially in fp32 -- the reference's own loop;
//   * anything that contradicts a guess falls back to that serial loop over the chunk or the
//     sub-chunk (slower, exact).
// One warp walks; the other warps of the CTA only prefetch and, at the end, fill in the boundary
// values of the chunks that were crossed in runs.
constexpr int kSpineNT = 512;
constexpr int kMaxSlots = 32;    // packs prefetched into shared memory (the rest are fetched when reached)
constexpr int kStageWords = kSub + kSubsPerChunk * (kSub + 1) + 32;  // scratch line + one padded chunk
constexpr int kMaxEvents = 640;  // precompiled chunk-level steps (more: the general walk is used)

__host__ __device__ inline int spine_slots(int nchunks) {  // as many as fit beside the per-chunk arrays
    const long room = 214L * 1024 - ((long)(nchunks + 1) * 8 + (long)nchunks * 13 + (long)((nchunks + 32) / 32) * 4) -
                      (long)kStageWords * 4 - (long)kMaxEvents * 20;
    const long fit = room / (long)sizeof(Pack) - 1;
    return (int)(fit > kMaxSlots ? kMaxSlots : (fit < 4 ? 4 : fit));
}
__host__ __device__ inline size_t spine_smem_bytes(int nchunks) {
    size_t b = (size_t)(nchunks + 1) * 8 + (size_t)nchunks * 13 + (size_t)((nchunks + 32) / 32) * 4;
    b = (b + 127) & ~(size_t)127;
    return b + (size_t)(spine_slots(nchunks) + 1) * sizeof(Pack) + (size_t)kStageWords * 4 + (size_t)kMaxEvents * 20;
}

// exact pass over `count` consecutive candidates held in shared memory, from state b: this is
// the reference's own loop (`sum += val` in fp32), run redundantly by every lane of the warp.
// `v` must point into SHARED memory (16-byte aligned).  All loads are issued first (explicit
// ld.shared, so that they are neither generic nor throttled by the scheduler's register heuristics);
// the 64 dependent adds then run back to back, 4 cycles each.
__device__ __forceinline__ float serial_block(float s, uint32_t saddr) {
    uint4 q[kSub / 4];
#pragma unroll
    for (int i = 0; i < kSub / 4; ++i)
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q[i].x), "=r"(q[i].y), "=r"(q[i].z), "=r"(q[i].w) : "r"(saddr + 16u * i));
#pragma unroll
    for (int i = 0; i < kSub / 4; ++i) {
        s = __fadd_rn(s, b2f(q[i].x));
        s = __fadd_rn(s, b2f(q[i].y));
        s = __fadd_rn(s, b2f(q[i].z));
        s = __fadd_rn(s, b2f(q[i].w));
    }
    return s;
}
__device__ __forceinline__ uint32_t serial_chain(uint32_t b, const uint32_t* v, int count) {
    float s = b2f(b);
    const uint32_t saddr = smem_u32(v);
    for (int blk = 0; blk < count / kSub; ++blk) s = serial_block(s, saddr + blk * kSub * 4u);
    return f2b(s);
}

// lowest set bit position >= from in a bitmask stored as 32-bit words (a sentinel bit always exists)
__device__ __forceinline__ int next_bit(const uint32_t* words, int from) {
    int w = from >> 5;
    uint32_t m = words[w] & (0xffffffffu << (from & 31));
    while (m == 0u) m = words[++w];
    return (w << 5) + __ffs(m) - 1;
}

struct SpineCtx {
    WeightSrc src; int64_t n; float shift, totalw; int lane; int* dbg; long long* dbgt;
#ifndef MB_PACK_PROF
#define MB_PACK_PROF 0
#endif
    bool prof_; long long* pacc;  // developer aid (MERCURY_SPINE_DEBUG=9): cycle accounting of the production walk
    // shared memory
    const unsigned long long* pc; const uint32_t* ce; const uint32_t* qd0; const uint32_t* qd1; const uint32_t* stopw;
    const uint8_t* slot_of; const Pack* slots; Pack* spare; uint32_t* stage;
    // global
    const Summary* sums; const Pack* packs; uint32_t* start_bits;
};

// inclusive scan of one map per lane, in lane order.  Without exact ties every map is a plain
// integer increment (d0 == d1): an integer prefix sum then does (32 * 2^25 fits in 32 bits).
__device__ __forceinline__ Pair warp_scan_pair(Pair x, int lane) {
    if (__ballot_sync(0xffffffffu, x.d0 != x.d1) == 0u) {
        uint32_t v = x.d0;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, v, off);
            if (lane >= off) v += o;
        }
        v = sat(v);
        return Pair{v, v};
    }
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        Pair o = shfl_up_pair(x, off);
        if (lane >= off) x = compose(o, x);
    }
    return x;
}

// fallback over a whole chunk whose guess was contradicted (no usable summary): stage it in shared
// memory, then go phase by phase -- every lane reduces "its" 64-candidate sub-chunk to a map for
// the current phase, a warp scan finds the sub-chunk in which the phase changes, that one is
// summed serially (the reference loop), and the remaining sub-chunks are redone in the new phase.
template <bool DBG>
__device__ __noinline__ uint32_t serial_chunk(const SpineCtx& x, int c, uint32_t b) {
    if (DBG && x.lane == 0) x.dbg[1]++;
    const int lane = x.lane;
    const int64_t e0 = (int64_t)c * kChunk;
    // layout: stage[0..63] is an aligned scratch line; sub-chunk k lives at stage[64 + 65*k ..] (the odd
    // stride keeps the 32 lanes, each walking its own sub-chunk, on different banks)
    uint32_t* padded = x.stage + kSub;
#pragma unroll 8
    for (int i = lane; i < kChunk; i += 32)
        padded[(i >> 6) * (kSub + 1) + (i & (kSub - 1))] = e0 + i < x.n ? f2b(x.src.get(e0 + i, x.shift, x.totalw)) : 0u;
    __syncwarp();
    int first = 0;
    while (first < kSubsPerChunk) {
        const uint32_t E = phase_of(b), S = b - (E << 23);
        Pair m = identity_pair();
        if (lane >= first) {
            const uint32_t* v = padded + lane * (kSub + 1);
            for (int j = 0; j < kSub; ++j) m = compose(m, addend_pair_fast(v[j], E));
        }
        const Pair inc = warp_scan_pair(m, lane);
        Pair ex = shfl_up_pair(inc, 1);
        if (lane == 0) ex = identity_pair();
        const uint32_t s_in = apply(ex, S), s_out = apply(inc, S);
        const unsigned over = __ballot_sync(0xffffffffu, lane >= first && s_out >= kPhaseLimit);
        if (!over) { b = bits_of(E, __shfl_sync(0xffffffffu, s_out, 31)); break; }
        const int f = __ffs(over) - 1;
        x.stage[2 * lane] = padded[f * (kSub + 1) + 2 * lane];
        x.stage[2 * lane + 1] = padded[f * (kSub + 1) + 2 * lane + 1];
        __syncwarp();
        b = serial_chain(bits_of(E, __shfl_sync(0xffffffffu, s_in, f)), x.stage, kSub);
        __syncwarp();
        first = f + 1;
    }
    __syncwarp();
    return b;
}

// serial pass over sub-chunk k of chunk c: raw candidates from the pack if it carries them
template <bool DBG>
__device__ __forceinline__ uint32_t serial_sub(const SpineCtx& x, const Pack* P, int c, int k, uint32_t b, int raw_slot = -2) {
    const long long t0 = DBG ? clock64() : 0;
    int r = raw_slot;  // (the precompiled walk knows where the raw candidates are; -2: look them up)
    if (r == -2) {
        r = -1;
#pragma unroll
        for (int q = 0; q < kRawPerPack; ++q) if (P->rawsub[q] == k) r = q;
    }
    const uint32_t* raw;
    if (DBG && x.lane == 0) x.dbg[r >= 0 ? 2 : 3]++;
    if (r >= 0) raw = P->raw[r];
    else {
        const int64_t i0 = (int64_t)c * kChunk + k * kSub + 2 * x.lane;
        x.stage[2 * x.lane] = i0 < x.n ? f2b(x.src.get(i0, x.shift, x.totalw)) : 0u;
        x.stage[2 * x.lane + 1] = i0 + 1 < x.n ? f2b(x.src.get(i0 + 1, x.shift, x.totalw)) : 0u;
        __syncwarp();
        raw = x.stage;
    }
    b = serial_chain(b, raw, kSub);
    __syncwarp();
    if (DBG && x.lane == 0) x.dbgt[0] += clock64() - t0;
    return b;
}

// a flagged chunk from sub-chunk k0 on, the general way: whatever the phase turns out to be
template <bool DBG>
__device__ __noinline__ uint32_t resolve_flagged(const SpineCtx& x, const Pack* P, int c, uint32_t b, int k0) {
    const int lane = x.lane;
    const uint32_t ek_l = P->e[lane];
    const uint32_t g_l = ek_l & kPhaseMask;
    const unsigned any_m = __ballot_sync(0xffffffffu, g_l == kAnyPhase);
    const unsigned flag_m = __ballot_sync(0xffffffffu, (ek_l & kFlag) != 0u);
    const unsigned tie0_m = __ballot_sync(0xffffffffu, (ek_l & kTie0) != 0u), tie1_m = __ballot_sync(0xffffffffu, (ek_l & kTie1) != 0u);
    // guess changes against the previous non-zero sub-chunk (defensive: expected only at flagged ones)
    const uint32_t prev_g = __shfl_up_sync(0xffffffffu, g_l, 1);
    const unsigned chg_m = __ballot_sync(0xffffffffu, lane > 0 && g_l != kAnyPhase && prev_g != kAnyPhase && prev_g != g_l);
    int k = k0;
    while (k < kSubsPerChunk) {
        const uint32_t E = phase_of(b), S = b - (E << 23);
        if ((any_m >> k) & 1u) { ++k; continue; }                   // all zeros: nothing happens
        const uint32_t g = P->e[k] & kPhaseMask;
        const int w = E == g ? 0 : (E == g + 1u ? 1 : -1);
        if (w < 0) { b = serial_sub<DBG>(x, P, c, k, b); ++k; continue; }
        const unsigned tie_m = w ? tie1_m : tie0_m;
        if (((tie_m | flag_m) >> k) & 1u) {                          // exact ties inside, or a possible phase change: apply the 2-state map
            const Pair q = w ? P->q1[k] : P->q0[k];
            const uint32_t Sn = apply(q, S);
            if (Sn >= kPhaseLimit) b = serial_sub<DBG>(x, P, c, k, b); else b = bits_of(E, Sn);
            ++k;
            continue;
        }
        // additive run [k, s): up to the next sub-chunk that needs individual treatment
        const unsigned stops = (flag_m | tie_m | chg_m) & ~((2u << k) - 1u);
        const int s = stops ? __ffs(stops) - 1 : kSubsPerChunk;
        const uint32_t* ps = w ? P->ps1 : P->ps0;
        const uint32_t base = ps[k];
        const uint32_t Sn = S + (ps[s] - base);
        if (Sn < kPhaseLimit) { b = bits_of(E, Sn); k = s; continue; }
        // the phase changes inside the run after all: find the sub-chunk, sum it serially
        const bool over = lane >= k && lane < s && S + (ps[lane + 1] - base) >= kPhaseLimit;
        const int f = __ffs(__ballot_sync(0xffffffffu, over)) - 1;
        b = serial_sub<DBG>(x, P, c, f, bits_of(E, S + (ps[f] - base)));
        k = f + 1;
    }
    return b;
}

// a flagged chunk along its precompiled walk (the expected case); the first contradicted guard
// hands the rest of the chunk to the general code
template <bool DBG>
__device__ __forceinline__ uint32_t run_pack(const SpineCtx& x, const Pack* P, int c, uint32_t b) {
    const int nev = (int)P->nev;
    uint32_t E = phase_of(b), S = b - (E << 23);
    long long tq = (MB_PACK_PROF && x.prof_) ? clock64() : 0;
    for (int i = 0; i < nev; ++i) {
        const uint4 ev = P->ev[i];  // (loading the next step ahead of time was measured: 1.2 us slower per walk)
        const uint32_t kind = ev.x & 0xffu, ph = (ev.x >> 8) & kPhaseMask;
        if (kind == kEvAdd) {
            const uint32_t Sn = S + ev.y;
            if ((ph == E || ph == kAnyPhase) && Sn < kPhaseLimit) { S = Sn; continue; }
        } else if (kind == kEvMap) {
            const uint32_t Sn = S + ((S & 1u) ? ev.w : ev.z);
            if (ph == E && Sn < kPhaseLimit) { S = Sn; continue; }
        } else {
            // (the map of a sub-chunk is exact iff the sum stays in the map's phase, i.e. iff S ends below 2^24)
            if (ph == E && !(ev.y & 0x100u) && S + ev.z < kPhaseLimit) { S += ev.z; continue; }
            if (ph + 1u == E && !(ev.y & 0x200u) && S + ev.w < kPhaseLimit) { S += ev.w; continue; }
            if (MB_PACK_PROF && x.prof_) { const long long t = clock64(); x.pacc[0] += t - tq; tq = t; }
            b = serial_sub<DBG>(x, P, c, (int)(ev.x >> 20), bits_of(E, S), (ev.y & 0xffu) == 0xffu ? -1 : (int)(ev.y & 0xffu));
            E = phase_of(b); S = b - (E << 23);
            if (MB_PACK_PROF && x.prof_) { const long long t = clock64(); x.pacc[2] += t - tq; x.pacc[3]++; tq = t; }
            continue;
        }
        if (DBG && x.lane == 0) x.dbg[7]++;
        if (MB_PACK_PROF && x.prof_) { const long long t = clock64(); x.pacc[0] += t - tq; tq = t; }
        const uint32_t rb = resolve_flagged<DBG>(x, P, c, bits_of(E, S), (int)(ev.x >> 20));
        if (MB_PACK_PROF && x.prof_) { x.pacc[4] += clock64() - tq; x.pacc[5]++; }
        return rb;
    }
    if (MB_PACK_PROF && x.prof_) { x.pacc[0] += clock64() - tq; x.pacc[1] += nev; }
    return bits_of(E, S);
}

// a pack that was not prefetched: one cooperative copy instead of many dependent global reads
template <bool DBG>
__device__ __noinline__ const Pack* fetch_pack(const SpineCtx& x, int c) {
    const uint4* g4 = reinterpret_cast<const uint4*>(x.packs + x.sums[c].pad[0]);
    uint4* s4 = reinterpret_cast<uint4*>(x.spare);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(Pack) / 16) / 32; ++i) s4[x.lane + 32 * i] = g4[x.lane + 32 * i];
    __syncwarp();
    return x.spare;
}
template <bool DBG>
__device__ __forceinline__ const Pack* pack_of(const SpineCtx& x, int c) {
    const int sl = (int)x.slot_of[c] - 1;
    if (DBG && x.lane == 0) { x.dbg[0]++; if (sl < 0) x.dbg[5]++; }
    return sl >= 0 ? &x.slots[sl] : fetch_pack<DBG>(x, c);
}

// chunks [c, c_end) the general way (no precompiled steps, or one of them was contradicted);
// writes their boundary values itself
template <bool DBG>
__device__ __noinline__ uint32_t walk_general(const SpineCtx& x, int c, int c_end, uint32_t b) {
    const int lane = x.lane;
    while (c < c_end) {
        const uint32_t E = phase_of(b), S = b - (E << 23);
        const uint32_t ec = x.ce[c], g = ec & kPhaseMask;
        if (lane == 0) x.start_bits[c] = b;
        if (ec & kFlag) {
            const Pack* P = pack_of<DBG>(x, c);
            b = run_pack<DBG>(x, P, c, b);
            __syncwarp();
            ++c;
            continue;
        }
        if (g == kAnyPhase) { ++c; continue; }  // all zeros
        if (g != E || (ec & kNoPack)) { b = serial_chunk<DBG>(x, c, b); ++c; continue; }
        if (ec & kTie0) {
            const uint32_t Sn = apply(Pair{x.qd0[c], x.qd1[c]}, S);
            if (Sn >= kPhaseLimit) b = serial_chunk<DBG>(x, c, b); else b = bits_of(E, Sn);
            ++c;
            continue;
        }
        // additive run [c, s): one subtraction of prefix sums
        const int s = min(next_bit(x.stopw, c + 1), c_end);
        const unsigned long long base = x.pc[c];
        const unsigned long long Sn = (unsigned long long)S + (x.pc[s] - base);
        int end = s;
        if (Sn >= kPhaseLimit) {  // the phase changes inside the run after all: first chunk that leaves it
            int lo = c, hi = s;   // state before chunk lo is inside the phase, before chunk hi outside
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if ((unsigned long long)S + (x.pc[mid] - base) >= kPhaseLimit) hi = mid; else lo = mid;
            }
            end = lo;
        }
        for (int k = c + 1 + lane; k < end; k += 32) x.start_bits[k] = bits_of(E, S + (uint32_t)(x.pc[k] - base));
        if (end == s) { b = bits_of(E, (uint32_t)Sn); c = s; continue; }
        const uint32_t bf = bits_of(E, S + (uint32_t)(x.pc[end] - base));
        if (lane == 0) x.start_bits[end] = bf;
        b = serial_chunk<DBG>(x, end, bf);
        c = end + 1;
    }
    return b;
}

// developer aid (MERCURY_SPINE_DEBUG=2): the precompiled walk through a pack against the general code
template <bool DBG>
__device__ __noinline__ uint32_t debug_check_pack(const SpineCtx& x, const Pack* P, int c, uint32_t b_in, uint32_t b_fast) {
    const uint32_t b_ref = resolve_flagged<DBG>(x, P, c, b_in, 0);
    if (b_ref != b_fast && x.lane == 0) {
        printf("PACK MISMATCH chunk %d: in=%08x fast=%08x general=%08x nev=%u\n", c, b_in, b_fast, b_ref, P->nev);
        for (int q = 0; q < (int)P->nev; ++q) printf("   ev[%d] = %08x %08x %08x %08x\n", q, P->ev[q].x, P->ev[q].y, P->ev[q].z, P->ev[q].w);
    }
    return b_ref;
}

template <bool DBG, bool PEER = false>
__global__ void __launch_bounds__(kSpineNT) spine_kernel(WeightSrc src, int64_t n, int nchunks, const Summary* __restrict__ sums,
                                                         const Pack* __restrict__ packs, uint32_t* __restrict__ start_bits,
                                                         float* __restrict__ total_out, int32_t* status, float* ema_out, float* chain_total_out, int debug_level,
                                                         PeerBox* const* peers, int nshards, int rank) {
    const int debug = DBG ? debug_level : 0;
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ int nslots_s, nev_s;
    __shared__ int slot_global[kMaxSlots];
    __shared__ unsigned long long wtot[kSpineNT / 32];
    __shared__ unsigned long long carry_s;
    __shared__ int wcnt[kSpineNT / 32];
    __shared__ int carry_ev;
    __shared__ long long dbg_t[4];
    __shared__ long long dbg_t2[8];  // cycles in serial passes, in packs (total), fetching packs
    __shared__ int dbg_n[8];  // flagged, serial chunks, serial subs (raw at hand), serial subs (fetched), events, late packs, general walks
    const bool prof = !DBG && debug_level == 9;   // developer aid: phase timestamps from the production kernel
    long long pt0 = 0, pt1 = 0, pt2 = 0, ptw = 0;
    if (prof && threadIdx.x == 0) ptw = clock64();
    pdl_wait();
    pdl_launch_dependents();
    if (prof && threadIdx.x == 0) pt0 = clock64();
    // the statistics the walk may need (fallback paths) and the EMA it publishes: loaded now, used microseconds later
    float stat_shift, stat_total;
    src.prepare(stat_shift, stat_total);
    const float stat_ema = (ema_out && src.stats) ? src.stats[MERCURY_STAT_EMA] : 0.f;
    // sharded pool with the exchange inside the chain: this kernel publishes the shard's chain total to every peer
    PeerBox* const my_box = PEER ? peers[rank] : nullptr;
    const unsigned int total_tag = PEER ? my_box->seq_total + 1u : 0u, sum_seq = PEER ? my_box->seq_sum : 0u;
    if (DBG && debug && threadIdx.x == 0) { dbg_t[0] = clock64(); for (int i = 0; i < 8; ++i) dbg_n[i] = 0; for (int i = 0; i < 4; ++i) dbg_t2[i] = 0; }
    // shared-memory layout
    unsigned long long* pc = reinterpret_cast<unsigned long long*>(sm);               // [nchunks+1] prefix of increments
    uint32_t* ce = reinterpret_cast<uint32_t*>(pc + nchunks + 1);                      // [nchunks] guess | flags
    uint32_t* qd0 = ce + nchunks; uint32_t* qd1 = qd0 + nchunks;                      // [nchunks] map for the guess
    uint32_t* stopw = qd1 + nchunks;                                                   // [(nchunks+32)/32] stop bitmask
    uint8_t* slot_of = reinterpret_cast<uint8_t*>(stopw + (nchunks + 32) / 32);
    size_t off = (size_t)(nchunks + 1) * 8 + (size_t)nchunks * 13 + (size_t)((nchunks + 32) / 32) * 4;
    off = (off + 127) & ~(size_t)127;
    const int max_slots = spine_slots(nchunks);
    Pack* slots = reinterpret_cast<Pack*>(sm + off);  // [max_slots] prefetched + 1 spare for late fetches
    Pack* spare = slots + max_slots;
    uint32_t* stage = reinterpret_cast<uint32_t*>(sm + off + (size_t)(max_slots + 1) * sizeof(Pack));
    uint4* evs = reinterpret_cast<uint4*>(stage + kStageWords);                              // [kMaxEvents] precompiled steps
    uint32_t* evb = reinterpret_cast<uint32_t*>(evs + kMaxEvents);                      // [kMaxEvents] state at each step
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- prefetch (whole CTA): summaries -> shared memory, prefix sums of the increments
    if (tid == 0) { nslots_s = 0; carry_s = 0ull; carry_ev = 0; }
    __syncthreads();
    for (int c0 = 0; c0 < nchunks; c0 += kSpineNT) {
        const int c = c0 + tid;
        unsigned long long inc = 0ull;
        if (c < nchunks) {
            const Summary s = sums[c];
            ce[c] = s.e; qd0[c] = s.q[0].d0; qd1[c] = s.q[0].d1;
            uint8_t so = 0;
            if (s.e & kFlag) {
                const int sl = atomicAdd(&nslots_s, 1);
                if (sl < max_slots) { so = (uint8_t)(sl + 1); slot_global[sl] = (int)s.pad[0]; }
            }
            slot_of[c] = so;
            inc = s.q[0].d0;
        }
        const unsigned long long mine = inc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) wtot[warp] = inc;
        __syncthreads();
        unsigned long long pre = carry_s;
        for (int i = 0; i < warp; ++i) pre += wtot[i];
        if (c < nchunks) pc[c] = pre + inc - mine;
        __syncthreads();
        if (tid == kSpineNT - 1) carry_s = pre + inc;
        __syncthreads();
    }
    if (tid == 0) pc[nchunks] = carry_s;
    // packs of the flagged chunks (as many as fit)
    const int nsl = min(nslots_s, max_slots);
    constexpr int kPackVecs = (int)(sizeof(Pack) / 16);
    for (int i = tid; i < nsl * kPackVecs; i += kSpineNT) {
        const int sl = i / kPackVecs, w = i % kPackVecs;
        reinterpret_cast<uint4*>(&slots[sl])[w] = reinterpret_cast<const uint4*>(packs + slot_global[sl])[w];
    }
    // stop mask: chunks that need individual treatment, or where the guess changes
    for (int w0 = 0; w0 < (nchunks + 32) / 32; w0 += kSpineNT / 32) {
        const int c = (w0 + warp) * 32 + lane;
        bool stop = false;
        if (c < nchunks) {
            const uint32_t e = ce[c], g = e & kPhaseMask;
            stop = (e & (kFlag | kTie0 | kNoPack)) != 0u;
            if (c > 0) { const uint32_t gp = ce[c - 1] & kPhaseMask; if (g != kAnyPhase && gp != kAnyPhase && g != gp) stop = true; }
        }
        if (c == nchunks) stop = true;  // sentinel
        const unsigned bits = __ballot_sync(0xffffffffu, stop);
        if (lane == 0 && w0 + warp < (nchunks + 32) / 32) stopw[w0 + warp] = bits;
    }
    __syncthreads();
    // precompiled chunk-level walk
    for (int c0 = 0; c0 < nchunks; c0 += kSpineNT) {
        const int c = c0 + tid;
        bool start = false;
        // (a step starts at chunk 0, at every stop, and after every chunk that is a step of its own)
        if (c < nchunks) start = c == 0 || ((stopw[c >> 5] >> (c & 31)) & 1u) || (ce[c - 1] & (kFlag | kTie0 | kNoPack)) != 0u;
        const unsigned bal = __ballot_sync(0xffffffffu, start);
        if (lane == 0) wcnt[warp] = __popc(bal);
        __syncthreads();
        int idx = carry_ev + __popc(bal & ((1u << lane) - 1u));
        for (int i = 0; i < warp; ++i) idx += wcnt[i];
        if (start && idx < kMaxEvents) {
            const uint32_t e = ce[c];
            uint4 ev;
            if (e & kFlag) ev = make_uint4(kEvPack, (uint32_t)c | ((uint32_t)(c + 1) << 16), 0u, 0u);
            else if (e & kNoPack) ev = make_uint4(kEvSlow, (uint32_t)c | ((uint32_t)(c + 1) << 16), 0u, 0u);
            else if (e & kTie0) ev = make_uint4(kEvMap | ((e & kPhaseMask) << 8), (uint32_t)c | ((uint32_t)(c + 1) << 16), qd0[c], qd1[c]);
            else {
                const int s = next_bit(stopw, c + 1);
                uint32_t g = kAnyPhase;
                for (int k = c; k < s && g == kAnyPhase; ++k) g = ce[k] & kPhaseMask;  // first chunk of the run with addends
                const unsigned long long inc = pc[s] - pc[c];
                ev = make_uint4(kEvAdd | (g << 8), (uint32_t)c | ((uint32_t)s << 16), (uint32_t)inc, (uint32_t)(inc >> 32));
            }
            evs[idx] = ev;
        }
        __syncthreads();
        if (tid == kSpineNT - 1) carry_ev = idx + (start ? 1 : 0);
        __syncthreads();
    }
    if (tid == 0) nev_s = carry_ev;
    __syncthreads();
    const int nev = nev_s;
    const bool use_events = nev <= kMaxEvents && nchunks < 65536;

    if (DBG && debug && tid == 0) dbg_t[1] = clock64();
    if (prof && tid == 0) pt1 = clock64();
    if (warp == 0) {
        // ---- the walk (one warp; scalar logic is replicated on all lanes)
        SpineCtx x;
        x.dbg = dbg_n;
        x.dbgt = dbg_t2;
        x.prof_ = prof; x.pacc = dbg_t2 + 0;
        if (prof && lane == 0) for (int i = 0; i < 8; ++i) dbg_t2[i] = 0;
        __syncwarp();
        x.src = src; x.n = n; x.lane = lane;
        x.shift = stat_shift; x.totalw = stat_total;
        x.pc = pc; x.ce = ce; x.qd0 = qd0; x.qd1 = qd1; x.stopw = stopw; x.slot_of = slot_of; x.slots = slots; x.spare = spare;
        x.stage = stage; x.sums = sums; x.packs = packs; x.start_bits = start_bits;
        uint32_t b = 0u;  // the running sum starts at +0
        if (DBG && debug == 3) {  // micro-timing of the primitives
            uint32_t t = 1u;
            long long c0 = clock64();
            for (int r = 0; r < 8; ++r) t = serial_chain(t, stage, kSub);
            long long c1 = clock64();
            uint32_t acc = t;
            for (int r = 0; r < 64; ++r) { const uint4 ev = evs[r & 15]; acc += ev.x + ev.y; if (acc & 1u) acc ^= ev.z; }
            long long c2 = clock64();
            if (lane == 0) printf("micro: serial_chain(64) = %lld cycles each; 64 dependent smem loads+alu = %lld cycles each (%u)\n", (c1 - c0) / 8, (c2 - c1) / 64, acc);
        }
        if (!use_events) {
            b = walk_general<DBG>(x, 0, nchunks, b);
        } else {
            uint32_t E = 0u, S = 0u;
            long long ppack = 0;
            int npack = 0, nsub = 0;
            for (int i = 0; i < nev; ++i) {
                const uint4 ev = evs[i];
                const uint32_t kind = ev.x & 0xffu, ph = (ev.x >> 8) & kPhaseMask;
                const int c = (int)(ev.y & 0xffffu), s = (int)(ev.y >> 16);
                if (kind == kEvAdd) {
                    const unsigned long long Sn = (unsigned long long)S + (((unsigned long long)ev.w << 32) | ev.z);
                    if ((ph == E || ph == kAnyPhase) && Sn < kPhaseLimit) {
                        if (lane == 0) evb[i] = bits_of(E, S);  // boundary values of [c, s): filled in by the whole CTA afterwards
                        S = (uint32_t)Sn;
                        continue;
                    }
                } else if (kind == kEvMap) {
                    const uint32_t Sn = S + ((S & 1u) ? ev.w : ev.z);
                    if (ph == E && Sn < kPhaseLimit) {
                        if (lane == 0) start_bits[c] = bits_of(E, S);
                        S = Sn;
                        continue;
                    }
                } else if (kind == kEvPack) {
                    b = bits_of(E, S);
                    if (lane == 0) start_bits[c] = b;
                    const long long tp0 = (DBG || prof) ? clock64() : 0;
                    const int sl = (int)slot_of[c] - 1;
                    if (DBG && debug && lane == 0) { dbg_n[0]++; if (sl < 0) dbg_n[5]++; }
                    if (sl < 0) fetch_pack<DBG>(x, c);
                    const Pack* P = sl >= 0 ? &slots[sl] : spare;  // (both in shared memory: plain LDS below)
                    if (DBG && debug && lane == 0) dbg_t2[2] += clock64() - tp0;
                    const uint32_t b_in = b;
                    b = run_pack<DBG>(x, P, c, b);
                    __syncwarp();
                    if (DBG && debug > 1) b = debug_check_pack<DBG>(x, P, c, b_in, b);
                    E = phase_of(b); S = b - (E << 23);
                    if (DBG && debug && lane == 0) dbg_t2[1] += clock64() - tp0;
                    if (prof) { ppack += clock64() - tp0; npack += 1; nsub += (int)P->nev; }
                    continue;
                }
                if (DBG && debug > 1 && lane == 0) printf("GENERAL step %d kind=%u ph=%u E=%u S=%u c=%d s=%d z=%u w=%u\n", i, kind, ph, E, S, c, s, ev.z, ev.w);
                // a contradicted guard, or a chunk without a shortcut: the general walk over this step's chunks
                if (DBG && lane == 0) x.dbg[6]++;
                if (lane == 0) evs[i].x = kEvSlow;  // (so that the fill below skips it)
                b = walk_general<DBG>(x, c, s, bits_of(E, S));
                E = phase_of(b); S = b - (E << 23);
            }
            b = bits_of(E, S);
            if (prof && lane == 0)
                printf("spine(prod): %d packs took %lld cycles, %d pack events | events %lld cyc | serial %lld cyc in %lld | fallbacks %lld cyc in %lld\n",
                       npack, ppack, nsub, dbg_t2[0], dbg_t2[2], dbg_t2[3], dbg_t2[4], dbg_t2[5]);
        }
        if (DBG && debug && lane == 0) { dbg_t[2] = clock64(); dbg_n[4] = nev; }
        if (prof && tid == 0) pt2 = clock64();
        if (lane == 0) {
            if (ema_out) *ema_out = stat_ema;  // fused statistics: EMAverage.value for the next call
            start_bits[nchunks] = b;
            const float total = b2f(b);
            *total_out = total;
            if (chain_total_out) *chain_total_out = total;
            if (!(total > 0.f)) atomicOr(status, MERCURY_ST_ZERO_SUM);
            else if (isinf(total)) atomicOr(status, MERCURY_ST_NONFINITE);
        }
        if constexpr (PEER) {
            for (int g = lane; g < nshards; g += 32)
                peer_store(&peers[g]->total[total_tag & 1u][rank], ((unsigned long long)b << 32) | total_tag);
            if (lane == 0) { my_box->seq_total = total_tag; my_box->seq_sum = sum_seq + 1u; }  // (read by the draw / the next step)
        }
    }
    __syncthreads();
    // ---- boundary values inside the additive steps (whole CTA, one warp per step)
    if (use_events) {
        for (int i = warp; i < nev; i += kSpineNT / 32) {
            const uint4 ev = evs[i];
            if ((ev.x & 0xffu) != kEvAdd) continue;
            const int c = (int)(ev.y & 0xffffu), s = (int)(ev.y >> 16);
            const uint32_t b0 = evb[i], E = phase_of(b0), S = b0 - (E << 23);
            const unsigned long long base = pc[c];
            for (int k = c + lane; k < s; k += 32) start_bits[k] = bits_of(E, S + (uint32_t)(pc[k] - base));
        }
    }
    if (prof) {
        __syncthreads();
        if (tid == 0) printf("spine(prod): wait=%lld prefetch=%lld walk=%lld fill=%lld cycles, events=%d\n", pt0 - ptw, pt1 - pt0, pt2 - pt1,
                             (long long)clock64() - pt2, nev);
    }
    if (DBG && debug) {
        __syncthreads();
        if (tid == 0)
            printf("spine: nchunks=%d prefetch=%lld walk=%lld fill=%lld cycles | flagged=%d (late packs %d) serial_chunks=%d "
                   "serial_subs=%d (+%d fetched) events=%d general_walks=%d pack_fallbacks=%d slots_wanted=%d | cycles: serial=%lld packs=%lld pack_of=%lld\n",
                   nchunks, dbg_t[1] - dbg_t[0], dbg_t[2] - dbg_t[1], (long long)clock64() - dbg_t[2], dbg_n[0], dbg_n[5], dbg_n[1],
                   dbg_n[2], dbg_n[3], dbg_n[4], dbg_n[6], dbg_n[7], nslots_s, dbg_t2[0], dbg_t2[1], dbg_t2[2]);
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
    int64_t n_weight;    // the N of N*p_i: the whole pool's size (== n unless the pool is sharded)
    int64_t idx_base;    // added to the positions written to idx[] (global position of this shard's first candidate)
    WeightSrc src; int64_t n; int nchunks;
    const uint32_t* start_bits; const float* total_ptr;
    // sharded pool with the exchange inside the chain: `u` are the GLOBAL uniforms; every CTA waits for the G chain totals
    // in its own box, picks the shard of its uniform and goes on only if that is this rank (shard_select_kernel's arithmetic)
    PeerBox* const* peers; int nshards, rank; float* totals_all; double* local_u; int32_t* status;
    const double* u; int64_t b;
    const unsigned char* data; int64_t row_bytes; const int64_t* labels; const int64_t* index;
    int64_t* idx; float* nprob; float* seed; unsigned char* out_data; int64_t* out_labels; int64_t* out_index;
    const double* approx;      // fp64 mode: running total of p in front of every chunk, [nchunks + 1]
    unsigned long long* dbg;   // developer aid: when set, CTA 0 stamps %globaltimer after the chunk's scores arrived and after the slot is known
};

// (d) of a draw: the payload row of candidate `pick` -> output row j, plus the scalar outputs.  Everybody issues its first
// loads at once; the scalar outputs ride along.
__device__ __forceinline__ void draw_gather(const DrawArgs& a, int64_t j, int64_t pick) {
    if (a.dbg && blockIdx.x == 0 && threadIdx.x == 64) a.dbg[1] = global_ns();
    const bool copy = a.data && a.out_data;
    const unsigned char* srow = copy ? a.data + pick * a.row_bytes : nullptr;
    unsigned char* drow = copy ? a.out_data + j * a.row_bytes : nullptr;
    const bool vec = copy && ((reinterpret_cast<uintptr_t>(srow) | reinterpret_cast<uintptr_t>(drow) | (uintptr_t)a.row_bytes) & 15) == 0;
    const int64_t nvec = vec ? (a.row_bytes >> 4) : 0;
    constexpr int kAhead = 4;
    uint4 v[kAhead];
#pragma unroll
    for (int k = 0; k < kAhead; ++k) {
        const int64_t i = threadIdx.x + (int64_t)k * kDrawNT;
        if (i < nvec) v[k] = ldg_stream(reinterpret_cast<const uint4*>(srow) + i);
    }
    if (threadIdx.x == 0) {
        // all loads first (a store in between would order them one latency after the other)
        const bool want_l = a.labels && a.out_labels, want_i = a.index && a.out_index;
        const int64_t lab = want_l ? a.labels[pick] : 0, ind = want_i ? a.index[pick] : 0;
        float shift, tot2;
        a.src.prepare(shift, tot2);
        const float p = a.src.get(pick, shift, tot2);
        // pytorch_collab.py:116: losses_pros[idx] * N ; :137,:145: grad = (1/B) / probs
        const float np_ = __fmul_rn(p, (float)a.n_weight);
        a.idx[j] = a.idx_base + pick;
        a.nprob[j] = np_;
        a.seed[j] = __fdiv_rn(__fdiv_rn(1.f, (float)a.b), np_);
        if (want_l) a.out_labels[j] = lab;
        if (want_i) a.out_index[j] = ind;
    }
    if (vec) {
#pragma unroll
        for (int k = 0; k < kAhead; ++k) {
            const int64_t i = threadIdx.x + (int64_t)k * kDrawNT;
            if (i < nvec) reinterpret_cast<uint4*>(drow)[i] = v[k];
        }
        for (int64_t i = threadIdx.x + (int64_t)kAhead * kDrawNT; i < nvec; i += kDrawNT)
            reinterpret_cast<uint4*>(drow)[i] = reinterpret_cast<const uint4*>(srow)[i];
    } else if (copy) {
        for (int64_t i = threadIdx.x; i < a.row_bytes; i += kDrawNT) drow[i] = srow[i];
    }
}

// barrier among the kDrawNT threads that work on a draw: the whole CTA in the stand-alone kernel, the 256 worker
// threads (named barrier 1) inside the fused step, whose CTAs carry one more warp
template <bool NAMED>
__device__ __forceinline__ void draw_sync() {
    if constexpr (NAMED) asm volatile("bar.sync 1, 256;" ::: "memory");
    else __syncthreads();
}

struct DrawSmem {
    ResolveSmem rs;
    int wcnt[kDrawNT / 32];
    int64_t chosen;
    double dwt[kDrawNT / 32];
};

// One draw by kDrawNT threads: `chunk` < 0 -> find the chunk from the boundary values first.  cdf_s: kChunk floats of
// shared memory.  `bits_known`: the exact running sum in front of `chunk` when the caller already has it.
template <bool NAMED>
__device__ __forceinline__ void draw_resolve(const DrawArgs& a, int64_t j, double uj, int chunk, uint32_t bits_known, bool have_bits,
                                             DrawSmem& ds, float* cdf_s, bool preloaded = false);

template <bool PEER>
__global__ void __launch_bounds__(kDrawNT) draw_kernel(DrawArgs a) {
    __shared__ DrawSmem ds;
    __shared__ float cdf_s[kChunk];
    pdl_wait();
    pdl_launch_dependents();
    const int64_t j = blockIdx.x;
    double uj = a.u[j];
    if constexpr (PEER) {
        __shared__ float sh_tot[kMaxShards];
        __shared__ double sh_u;
        __shared__ int sh_bad;
        PeerBox* mine = a.peers[a.rank];
        const unsigned int tag = mine->seq_total;   // (set by this step's spine)
        const int g = threadIdx.x;
        if (g == 0) sh_bad = (*a.status & MERCURY_ST_PEER_TIMEOUT) ? 1 : 0;   // (the sums of this step may have timed out already)
        __syncthreads();
        if (g < a.nshards) {
            unsigned long long w;
            if (!peer_wait(&mine->total[tag & 1u][g], tag, w, mine)) { atomicOr(a.status, MERCURY_ST_PEER_TIMEOUT); sh_bad = 1; }
            sh_tot[g] = __uint_as_float((unsigned int)(w >> 32));
            if (j == 0) a.totals_all[g] = sh_tot[g];
        }
        __syncthreads();
        if (threadIdx.x == 0) {   // serial fp32 CDF over the shard totals, searched as ATen does (as shard_select_kernel)
            float cs = 0.f;
            for (int k = 0; k < a.nshards; ++k) cs = __fadd_rn(cs, sh_tot[k]);
            float run = 0.f, lo_f = 0.f, hi_f = 0.f;
            int k = 0;
            for (;; ++k) {
                lo_f = __fdiv_rn(run, cs);            // (cdf of the previous shard: run / total, 0 for the first)
                run = __fadd_rn(run, sh_tot[k]);
                hi_f = __fdiv_rn(run, cs);
                if (k >= a.nshards - 1 || !((double)hi_f < uj)) break;
            }
            double out = -1.0;
            if (k == a.rank) {
                const double lo = k > 0 ? (double)lo_f : 0.0, hi = (double)hi_f;
                out = hi > lo ? (uj - lo) / (hi - lo) : 0.0;
                if (out < 0.0) out = 0.0;
                if (out >= 1.0) out = 0.99999999999999989;
            }
            if (sh_bad) out = peer_poison();
            sh_u = out;
            a.local_u[j] = out;
        }
        __syncthreads();
        uj = sh_u;
    }
    draw_resolve<false>(a, j, uj, -1, 0u, false, ds, cdf_s);
}

template <bool NAMED>
__device__ __forceinline__ void draw_resolve(const DrawArgs& a, int64_t j, double uj, int chunk, uint32_t bits_known, bool have_bits,
                                             DrawSmem& ds, float* cdf_s, bool preloaded) {
    ResolveSmem& rs = ds.rs;
    int* wcnt = ds.wcnt;
    int64_t& chosen = ds.chosen;
    const float total = *a.total_ptr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (!(uj >= 0.0)) {  // sharded pools: this draw belongs to another shard -> all-zero outputs (summed across ranks later);
        if (threadIdx.x == 0) {   // NaN: a peer exchange of this step timed out -> poisoned outputs
            const bool bad = uj != uj;
            const float q = bad ? __int_as_float(0x7fc00000) : 0.f;
            a.idx[j] = bad ? -1 : 0; a.nprob[j] = q; a.seed[j] = q;
            if (a.labels && a.out_labels) a.out_labels[j] = 0;
            if (a.index && a.out_index) a.out_index[j] = 0;
        }
        if (a.data && a.out_data) {
            unsigned char* drow = a.out_data + j * a.row_bytes;
            for (int64_t i = threadIdx.x; i < a.row_bytes; i += kDrawNT) drow[i] = 0;
        }
        return;
    }
    // (a) chunks whose last cdf value is < u
    if (chunk < 0) {
        int cnt = 0;
        for (int c = threadIdx.x; c < a.nchunks; c += kDrawNT) {
            const float endv = __fdiv_rn(b2f(a.start_bits[c + 1]), total);
            cnt += ((double)endv < uj) ? 1 : 0;
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) wcnt[warp] = cnt;
        draw_sync<NAMED>();
        chunk = 0;
#pragma unroll
        for (int i = 0; i < kDrawNT / 32; ++i) chunk += wcnt[i];
        if (chunk > a.nchunks - 1) chunk = a.nchunks - 1;
        draw_sync<NAMED>();
    }
    // (b) the slot inside that chunk, sub-chunk by sub-chunk (64 candidates, 8 consecutive threads each).  Per round:
    //     every sub-chunk is reduced to its map for the current phase; warp 0 scans the 32 maps from the sub-chunk at
    //     which the exact sum is known, which gives the exact sum after every sub-chunk up to the one where the phase
    //     changes; if u falls before that one, its sub-chunk is walked with the reference's own loop and we are done;
    //     otherwise the sub-chunk with the phase change is walked that way and the round repeats behind it.  A chunk
    //     in one phase takes one round; the head of the pool, where the sum doubles every few candidates, a handful
    //     (all the early changes fall inside the first sub-chunks and cost one serial walk each, not one round each).
    const int64_t start = (int64_t)chunk * kChunk;
    const int len = (int)min((int64_t)kChunk, a.n - start);
    uint32_t pb[kDrawIPT];
    float* pv = cdf_s;                       // the chunk's probabilities, for the serial walks
    if (preloaded) {   // (block-uniform: the caller put this very chunk's probabilities into cdf_s ahead of time, same layout)
        const int base = threadIdx.x * kDrawIPT;
#pragma unroll
        for (int k = 0; k < kDrawIPT; ++k) pb[k] = f2b(pv[base + k]);
    } else {
        load_chunk<kDrawNT, kDrawIPT>(a.src, start, len, pb);
        const int base = threadIdx.x * kDrawIPT;
#pragma unroll
        for (int k = 0; k < kDrawIPT; ++k) pv[base + k] = b2f(pb[k]);
    }
    uint32_t bits = have_bits ? bits_known : a.start_bits[chunk];
    Pair* subm = rs.tot;                     // [32] one map per sub-chunk
    if (a.dbg && blockIdx.x == 0 && threadIdx.x == 64) a.dbg[0] = global_ns();
    int cur = 0;                             // the exact sum `bits` is the one before sub-chunk `cur`
    for (;;) {                               // (block-uniform control flow: decisions travel through rs)
        const uint32_t E = phase_of(bits), S = bits - (E << 23);
        Pair m = identity_pair();
#pragma unroll
        for (int k = 0; k < kDrawIPT; ++k) m = compose(m, addend_pair_fast(pb[k], E));
#pragma unroll
        for (int off = 1; off < 8; off <<= 1) {
            Pair o;
            o.d0 = __shfl_down_sync(0xffffffffu, m.d0, off, 8); o.d1 = __shfl_down_sync(0xffffffffu, m.d1, off, 8);
            if ((lane & 7) + off < 8) m = compose(m, o);
        }
        if ((lane & 7) == 0) subm[threadIdx.x >> 3] = m;
        draw_sync<NAMED>();
        if (warp == 0) {
            const Pair mine = lane >= cur ? subm[lane] : identity_pair();
            const Pair inc = warp_scan_pair(mine, lane);
            const uint32_t s_out = apply(inc, S);                                 // state after sub-chunk `lane` ...
            const unsigned over = __ballot_sync(0xffffffffu, lane >= cur && s_out >= kPhaseLimit);
            const int f = over ? __ffs(over) - 1 : kSubsPerChunk;                 // ... exact for lane < f
            const bool reach = lane >= cur && lane < f && !((double)__fdiv_rn(b2f(bits_of(E, s_out)), total) < uj);
            const unsigned hit = __ballot_sync(0xffffffffu, reach);
            int target = -1;                 // sub-chunk to walk serially
            bool final_walk = false;
            if (hit) { target = __ffs(hit) - 1; final_walk = true; }
            else if (f < kSubsPerChunk) target = f;
            if (target < 0) {                // every sub-chunk ends below u (only the clamped last chunk): last slot
                if (lane == 0) { rs.cross = len - 1; rs.next_bits = 0u; rs.end_bits = 1u; }
            } else {
                const uint32_t s_prev = __shfl_sync(0xffffffffu, s_out, target > 0 ? target - 1 : 0);
                const uint32_t b_in = bits_of(E, target > cur ? s_prev : S);
                // the reference's loop over the 64 candidates; prefix values are kept for the count below
                float acc = b2f(b_in);
                const float* v = pv + target * kSub;
                float pre0 = 0.f, pre1 = 0.f;
#pragma unroll 16
                for (int i = 0; i < kSub; ++i) {
                    acc = __fadd_rn(acc, v[i]);
                    if (i == 2 * lane) pre0 = acc;
                    if (i == 2 * lane + 1) pre1 = acc;
                }
                const uint32_t b_out = f2b(acc);
                const int below = (((double)__fdiv_rn(pre0, total) < uj) ? 1 : 0) + (((double)__fdiv_rn(pre1, total) < uj) ? 1 : 0);
                const int cnt_in = __reduce_add_sync(0xffffffffu, below);
                if (final_walk || cnt_in < kSub) {           // u falls inside this sub-chunk
                    int t = target * kSub + cnt_in;
                    if (t > len - 1) t = len - 1;
                    if (lane == 0) { rs.cross = t; rs.end_bits = 1u; }
                } else if (target + 1 >= kSubsPerChunk) {    // walked the last sub-chunk and still below u: clamp
                    if (lane == 0) { rs.cross = len - 1; rs.end_bits = 1u; }
                } else {
                    if (lane == 0) { rs.cross = target + 1; rs.next_bits = b_out; rs.end_bits = 0u; }
                }
            }
        }
        draw_sync<NAMED>();
        const bool done = rs.end_bits != 0u;
        const int val = rs.cross;
        const uint32_t nb = rs.next_bits;
        draw_sync<NAMED>();
        if (done) {
            if (threadIdx.x == 0) chosen = start + val;
            break;
        }
        cur = val;
        bits = nb;
    }
    draw_sync<NAMED>();
    draw_gather(a, j, chosen);
}

// ---------------------------------------------------------------------------------------------------------------
// fp64 mode of the whole step (MERCURY_CDF_F64): cdf_i = float(fp64 running total of p / fp64 total), searched as ATen
// searches.  Deterministic, no serial walk -- and not index-identical to torch.multinomial (3.5 % of the
// draws agree at a million candidates); for callers that want a draw from the same distribution, not the same stream.
__global__ void __launch_bounds__(256) psum_kernel(WeightSrc src, int64_t n, double* __restrict__ sums, int32_t* status) {
    __shared__ double red[8];
    pdl_wait();
    pdl_launch_dependents();
    float shift, total;
    src.prepare(shift, total);
    const int64_t start = (int64_t)blockIdx.x * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    double acc = 0.0;
    int bad = 0;
    for (int i = threadIdx.x; i < len; i += 256) {
        const float x = src.get(start + i, shift, total);
        if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
        else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
        acc += (double)x;
    }
    if (bad) atomicOr(status, bad);
    const double t = block_sum<256>(acc, red);
    if (threadIdx.x == 0) sums[blockIdx.x] = t;
}

// One fp64 draw by kDrawNT threads.  chunk < 0: find it from a.approx first; else run0 is the running total in front of it.
template <bool NAMED>
__device__ __forceinline__ void draw_resolve_f64(const DrawArgs& a, int64_t j, double uj, int chunk, double run0, double total64,
                                                 DrawSmem& ds) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (chunk < 0) {
        int cnt = 0;
        for (int c = threadIdx.x; c < a.nchunks; c += kDrawNT) cnt += ((double)(float)(a.approx[c + 1] / total64) < uj) ? 1 : 0;
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) ds.wcnt[warp] = cnt;
        draw_sync<NAMED>();
        chunk = 0;
#pragma unroll
        for (int i = 0; i < kDrawNT / 32; ++i) chunk += ds.wcnt[i];
        if (chunk > a.nchunks - 1) chunk = a.nchunks - 1;
        run0 = a.approx[chunk];
        draw_sync<NAMED>();
    }
    const int64_t start = (int64_t)chunk * kChunk;
    const int len = (int)min((int64_t)kChunk, a.n - start);
    uint32_t pb[kDrawIPT];
    load_chunk<kDrawNT, kDrawIPT>(a.src, start, len, pb);
    double loc[kDrawIPT];
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kDrawIPT; ++k) { t += (double)b2f(pb[k]); loc[k] = t; }
    double inc = t;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const double o = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += o;
    }
    if (lane == 31) ds.dwt[warp] = inc;
    draw_sync<NAMED>();
    double pre = run0 + (inc - t);
    for (int i = 0; i < warp; ++i) pre += ds.dwt[i];
    // the slot: how many candidates of the chunk lie below u (the cdf never decreases)
    const int base = threadIdx.x * kDrawIPT;
    int below = 0;
#pragma unroll
    for (int k = 0; k < kDrawIPT; ++k)
        if (base + k < len && (double)(float)((pre + loc[k]) / total64) < uj) ++below;
    below = __reduce_add_sync(0xffffffffu, below);
    if (lane == 0) ds.wcnt[warp] = below;
    draw_sync<NAMED>();
    int slot = 0;
#pragma unroll
    for (int i = 0; i < kDrawNT / 32; ++i) slot += ds.wcnt[i];
    if (slot > len - 1) slot = len - 1;
    draw_sync<NAMED>();
    draw_gather(a, j, start + slot);
}

__global__ void __launch_bounds__(kDrawNT) draw_f64_kernel(DrawArgs a, float* chain_total_out, int32_t* status) {
    __shared__ DrawSmem ds;
    pdl_wait();
    pdl_launch_dependents();
    const double total64 = a.approx[a.nchunks];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (chain_total_out) *chain_total_out = (float)total64;
        if (!(total64 > 0.0)) atomicOr(status, MERCURY_ST_ZERO_SUM);
        else if (isinf((float)total64)) atomicOr(status, MERCURY_ST_NONFINITE);
    }
    draw_resolve_f64<false>(a, blockIdx.x, a.u[blockIdx.x], -1, 0.0, total64, ds);
}

#include "step_fused.cuh"

// workspace carving (all offsets 256-byte aligned)
struct CdfWs {
    double* sums; double* approx; int* slotid; Summary* summaries; Pack* subs; uint32_t* start_bits; float* total;
    StepCtrl* ctrl;  // whole-step layout only
    size_t bytes;
};
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
inline CdfWs carve(void* ws, int64_t n, bool direct_subs = false, int compact_slots = kSubSlots) {
    const size_t nchunks = (size_t)((n + kChunk - 1) / kChunk);
    const size_t nsubs = direct_subs && nchunks <= (size_t)kFusedMaxChunks ? (nchunks > (size_t)kSubSlots ? nchunks : (size_t)kSubSlots)
                                                                          : (size_t)compact_slots;
    // whole-step layout: one fp64 sum per K1 tile; the smallest tile is 64 rows (small_tile_rows: rows of 185..256 bytes).
    // (Until round 2 this said n / 128 -- right when the smallest tile had 128 rows; with 64-row tiles the sums of a pool of
    // wide rows overran into the approx / summaries regions, found by tools/fuzz_step.py as a spurious "negative" status.)
    const size_t nsums = direct_subs ? (size_t)(n / 64 + 2) : nchunks;
    unsigned char* p = static_cast<unsigned char*>(ws);
    CdfWs w;
    size_t off = 0;
    w.sums = reinterpret_cast<double*>(p + off); off += align256(nsums * 8);
    w.approx = reinterpret_cast<double*>(p + off); off += align256((nchunks + 1) * 8);
    w.summaries = reinterpret_cast<Summary*>(p + off); off += align256(nchunks * sizeof(Summary));
    w.subs = reinterpret_cast<Pack*>(p + off); off += align256(nsubs * sizeof(Pack));
    w.slotid = reinterpret_cast<int*>(p + off); off += align256(nchunks * 4);
    w.start_bits = reinterpret_cast<uint32_t*>(p + off); off += align256((nchunks + 1) * 4);
    w.total = reinterpret_cast<float*>(p + off); off += 256;
    w.ctrl = nullptr;
    if (direct_subs) { w.ctrl = reinterpret_cast<StepCtrl*>(p + off); off += align256(sizeof(StepCtrl)); }
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
int run_exact_boundaries(const WeightSrc& src, int64_t n, const CdfWs& w, int32_t* status, int check, cudaStream_t st,
                         const FusedStats* fused = nullptr, float* ema_out = nullptr) {
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    if (fused) {
        cudaError_t e = fused->peers
            ? launch_chain(true, summarize_kernel<true, true>, dim3(nchunks), dim3(kSumNT), 0, st, src, n, (const double*)w.approx,
                           (const int*)w.slotid, g_guess_bias.load(), w.summaries, w.subs, status, check, *fused)
            : launch_chain(true, summarize_kernel<true, false>, dim3(nchunks), dim3(kSumNT), 0, st, src, n, (const double*)w.approx,
                           (const int*)w.slotid, g_guess_bias.load(), w.summaries, w.subs, status, check, *fused);
        if (e != cudaSuccess) return cuda_fail(e, "summarize_kernel launch");
    } else {
        summarize_kernel<false, false><<<nchunks, kSumNT, 0, st>>>(src, n, w.approx, w.slotid, g_guess_bias.load(), w.summaries, w.subs, status,
                                                            check, FusedStats{});
    }
    MB_CHECK_LAUNCH("summarize_kernel");
    const size_t smem = spine_smem_bytes(nchunks);
    // opt in to the large dynamic shared memory, once per DEVICE and instantiation (a function's attributes belong to
    // one context: a second GPU of the same process needs its own call)
    {
        const size_t want = 226 * 1024;
        int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(spine_kernel<false, false>), want, "spine_kernel smem attribute");
        if (!rc) rc = ensure_dynamic_smem(reinterpret_cast<const void*>(spine_kernel<false, true>), want, "spine_kernel smem attribute");
        if (!rc) rc = ensure_dynamic_smem(reinterpret_cast<const void*>(spine_kernel<true, false>), want, "spine_kernel smem attribute");
        if (!rc) rc = ensure_dynamic_smem(reinterpret_cast<const void*>(spine_kernel<true, true>), want, "spine_kernel smem attribute");
        if (rc) return rc;
    }
    static const int debug = env_knob("MERCURY_SPINE_DEBUG", 0, 0, 9);  // developer aid (printf)
    float* chain_total_out = fused ? fused->stats + MERCURY_STAT_CHAIN_TOTAL : nullptr;
    PeerBox* const* peers = fused ? fused->peers : nullptr;
    const int nshards = fused ? fused->nshards : 0, rank = fused ? fused->rank : 0;
    cudaError_t e;
    if (debug && debug != 9) {
        e = peers ? launch_chain(false, spine_kernel<true, true>, dim3(1), dim3(kSpineNT), smem, st, src, n, nchunks, (const Summary*)w.summaries,
                                 (const Pack*)w.subs, w.start_bits, w.total, status, ema_out, chain_total_out, debug, peers, nshards, rank)
                  : launch_chain(false, spine_kernel<true, false>, dim3(1), dim3(kSpineNT), smem, st, src, n, nchunks, (const Summary*)w.summaries,
                                 (const Pack*)w.subs, w.start_bits, w.total, status, ema_out, chain_total_out, debug, peers, nshards, rank);
    } else {
        e = peers ? launch_chain(true, spine_kernel<false, true>, dim3(1), dim3(kSpineNT), smem, st, src, n, nchunks, (const Summary*)w.summaries,
                                 (const Pack*)w.subs, w.start_bits, w.total, status, ema_out, chain_total_out, debug, peers, nshards, rank)
                  : launch_chain(true, spine_kernel<false, false>, dim3(1), dim3(kSpineNT), smem, st, src, n, nchunks, (const Summary*)w.summaries,
                                 (const Pack*)w.subs, w.start_bits, w.total, status, ema_out, chain_total_out, debug, peers, nshards, rank);
    }
    if (e != cudaSuccess) return cuda_fail(e, "spine_kernel launch");
    MB_CHECK_LAUNCH("spine_kernel");
    return MERCURY_OK;
}

}  // namespace

size_t cdf_ws_bytes(int64_t n) { return carve(nullptr, n < 1 ? 1 : n).bytes; }
size_t update_ws_bytes(int64_t n) { return carve(nullptr, n < 1 ? 1 : n, true).bytes; }

int pool_stats_launch(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds, float* ema_state, int ema_first,
                      double ema_alpha, double smooth_alpha, float* stats, void* ws, cudaStream_t st, bool update_layout) {
    const CdfWs w = carve(ws, n, update_layout);
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
                                              stats, w.approx, w.slotid, g_guess_bias.load(), kSubSlots);
    MB_CHECK_LAUNCH("stats_finalize_kernel");
    return MERCURY_OK;
}

int sharded_step_launch(const mercury_update_args& a, const float* scores, const double* part, int per, void* ws, cudaStream_t st);

// statistics + boundaries + draw for the whole-step entry point.  `part`/`per` are the partial
// sums K1 produced on the way (per == 0: none, a chunk-sum pass runs instead).
int update_tail_launch(const mercury_update_args& a, const float* scores, const double* part, int per, void* ws, cudaStream_t st) {
    if (a.peer_boxes) return sharded_step_launch(a, scores, part, per, ws, st);
    const CdfWs w = carve(ws, a.n, true);
    const int nchunks = (int)((a.n + kChunk - 1) / kChunk);
    WeightSrc src{scores, a.stats};
    int rc = MERCURY_OK;
    if (a.cdf_mode == MERCURY_CDF_F64) {
        // statistics -> fp64 chunk sums of p -> their scan -> one CTA per draw (no serial walk)
        if (!a.draw_only) {
            rc = pool_stats_launch(scores, a.n, a.h_bounds, a.n_bounds, a.ema_state, a.ema_first, a.ema_alpha, a.smooth_alpha, a.stats,
                                   ws, st, true);
            if (rc) return rc;
            cudaError_t e = launch_chain(false, psum_kernel, dim3(nchunks), dim3(256), 0, st, src, a.n, w.sums, a.status);
            if (e != cudaSuccess) return cuda_fail(e, "psum_kernel launch");
            prefix_finalize_kernel<<<1, 1024, 0, st>>>(w.sums, nchunks, w.approx, w.slotid, 1.0, 0);
            MB_CHECK_LAUNCH("prefix_finalize_kernel");
        }
        if (a.b > 0) {
            DrawArgs d{};
            d.n_weight = a.n; d.idx_base = 0;
            d.src = src; d.n = a.n; d.nchunks = nchunks; d.start_bits = w.start_bits; d.total_ptr = w.total; d.approx = w.approx;
            d.u = a.uniforms; d.b = a.b;
            d.data = static_cast<const unsigned char*>(a.data); d.row_bytes = a.row_bytes; d.labels = a.labels; d.index = a.index;
            d.idx = a.idx; d.nprob = a.nprob; d.seed = a.seed; d.out_data = static_cast<unsigned char*>(a.out_data);
            d.out_labels = a.out_labels; d.out_index = a.out_index;
            cudaError_t e = launch_chain(false, draw_f64_kernel, dim3((unsigned)a.b), dim3(kDrawNT), 0, st, d,
                                         a.stats + MERCURY_STAT_CHAIN_TOTAL, a.status);
            if (e != cudaSuccess) return cuda_fail(e, "draw_f64_kernel launch");
        }
        return MERCURY_OK;
    }
    if (a.draw_only) {
        // boundaries, total and statistics are those a previous call on this workspace left behind
    } else if (a.n_bounds == 0 && nchunks <= kFusedMaxChunks) {
        if (per == 0) {
            chunk_sums_kernel<<<nchunks, 256, 0, st>>>(scores, a.n, w.sums, nullptr, 0);
            MB_CHECK_LAUNCH("chunk_sums_kernel");
            part = w.sums;
            per = kChunk;
        }
        FusedStats fs{};
        fs.part = part; fs.per = per; fs.npart = (int)((a.n + per - 1) / per);
        fs.ema_in = a.ema_state; fs.ema_first = a.ema_first; fs.ema_alpha = a.ema_alpha; fs.smooth_alpha = a.smooth_alpha;
        fs.stats = a.stats;
        fs.gsums = a.global_sums; fs.ngsums = a.n_global_sums; fs.n_global = a.global_sums ? a.n_global : a.n;
        rc = run_exact_boundaries(src, a.n, w, a.status, 1, st, &fs, a.ema_state);
    } else {
        rc = pool_stats_launch(scores, a.n, a.h_bounds, a.n_bounds, a.ema_state, a.ema_first, a.ema_alpha, a.smooth_alpha, a.stats,
                               ws, st, true);
        if (rc) return rc;
        rc = run_exact_boundaries(src, a.n, w, a.status, 1, st);
    }
    if (rc) return rc;
    if (a.b > 0) {
        DrawArgs d{};
        d.n_weight = a.global_sums ? a.n_global : a.n; d.idx_base = a.global_sums ? a.index_base : 0;
        d.src = src; d.n = a.n; d.nchunks = nchunks; d.start_bits = w.start_bits; d.total_ptr = w.total;
        d.u = a.uniforms; d.b = a.b;
        d.data = static_cast<const unsigned char*>(a.data); d.row_bytes = a.row_bytes; d.labels = a.labels; d.index = a.index;
        d.idx = a.idx; d.nprob = a.nprob; d.seed = a.seed; d.out_data = static_cast<unsigned char*>(a.out_data);
        d.out_labels = a.out_labels; d.out_index = a.out_index;
        cudaError_t e = launch_chain(true, draw_kernel<false>, dim3((unsigned)a.b), dim3(kDrawNT), 0, st, d);
        if (e != cudaSuccess) return cuda_fail(e, "draw_kernel launch");
    }
    return MERCURY_OK;
}


// ---- the whole step as one cooperative kernel (step_fused.cuh).  Returns 1 when the call does not fit that kernel (the
// caller then runs the four-kernel chain), MERCURY_OK when it was launched, a negative error otherwise.
std::atomic<int> g_fused_mode{-1};  // test hook (mercury_test_set_fused): -1 = environment (MERCURY_FUSED, default on), 0 = off, 1 = on

namespace {
std::mutex g_fused_mu;
std::map<std::pair<int, const void*>, int> g_fused_ctas;  // (device, kernel) -> co-resident CTAs per SM (0: cannot run)

template <typename T, int CT>
int fused_launch_typed(StepArgs& sa, int64_t want_ctas, cudaStream_t st) {
    auto kern = sa.cdf_mode == MERCURY_CDF_F64 ? fused_step_kernel<T, CT, true> : fused_step_kernel<T, CT, false>;
    const void* fn = reinterpret_cast<const void*>(kern);
    int rc = ensure_dynamic_smem(fn, kStepSmemBytes, "fused_step_kernel smem attribute");
    if (rc) return rc;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return 1;
    int per_sm;
    {
        std::lock_guard<std::mutex> lock(g_fused_mu);
        auto it = g_fused_ctas.find(std::make_pair(dev, fn));
        if (it == g_fused_ctas.end()) {
            int occ = 0, coop = 0;
            cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kStepThreads, kStepSmemBytes);
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
            const int v = (e == cudaSuccess && coop && occ > 0) ? (occ > 3 ? 3 : occ) : 0;
            it = g_fused_ctas.emplace(std::make_pair(dev, fn), v).first;
        }
        per_sm = it->second;
    }
    {   // developer knob: tile weights of the scoring phase
        static std::once_flag once;
        std::call_once(once, [] {
            const char* e = getenv("MERCURY_STEP_W");
            unsigned int w[5];
            if (e && sscanf(e, "%u,%u,%u,%u,%u", &w[0], &w[1], &w[2], &w[3], &w[4]) == 5 && w[0] && w[1] && w[2] && w[0] < 1000 && w[1] < 1000 && w[2] < 1000)
                cudaMemcpyToSymbol(c_tile_w, w, sizeof(w));
        });
    }
    const int sms = device_sm_count();
    if (per_sm <= 0 || sms <= 0) return 1;
    static const int knob_per_sm = env_knob("MERCURY_STEP_CTAS", 3, 1, 3);   // developer knob: CTAs per SM
    if (knob_per_sm < per_sm) per_sm = knob_per_sm;
    const int grid = (int)min_i64(min_i64((int64_t)per_sm * sms, want_ctas), kStepMaxCtas);
    {   // the split of the chunks over the worker CTAs (all but the last CTA), for the kernel's phase B
        const int nworkers = grid > 1 ? grid - 1 : 1;
        sa.chunks_q = sa.nchunks / nworkers; sa.chunks_r = sa.nchunks % nworkers; sa.parts_per_chunk = kChunk / sa.tile_rows;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kStepThreads); cfg.dynamicSmemBytes = kStepSmemBytes; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    // programmatic serialization behind the previous launch on the stream: the kernel's first instruction is the wait for
    // that launch, so only its dispatch and per-CTA set-up overlap the predecessor's tail
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    static const int step_pdl = env_knob("MERCURY_STEP_PDL", 1, 0, 1);
    cfg.attrs = attr; cfg.numAttrs = (step_pdl && pdl_enabled()) ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, sa);
    if (e != cudaSuccess) return cuda_fail(e, "fused_step_kernel launch");
    return MERCURY_OK;
}
}  // namespace

int fused_step_launch(const mercury_update_args& a, void* ws, cudaStream_t st) {
    static const int env_on = env_knob("MERCURY_FUSED", 1, 0, 1);
    const int fused_mode = g_fused_mode.load();
    if (!(fused_mode < 0 ? env_on : fused_mode)) return 1;
    if (a.scores_in || a.n_bounds != 0 || a.draw_only || a.global_sums) return 1;
    if (a.peer_boxes && a.cdf_mode != MERCURY_CDF_EXACT_SERIAL_F32) return 1;
    // sharded pools: the kernel chain is the faster one (measured at 2 and 8 GPUs: 54 us against 64 us per step -- two grid-wide
    // waits for the peers inside one resident grid cost more than inside two short kernels); the one-kernel sharded step stays
    // reachable through the test hook, which keeps it under test
    static const int peer_knob = env_knob("MERCURY_FUSED_PEER", 0, 0, 1);   // developer knob: the one-kernel step for sharded pools too
    if (a.peer_boxes && fused_mode != 1 && !peer_knob) return 1;

    if (a.b > kStepMaxDraws || a.row_stride != a.c) return 1;
    const int64_t elt = a.dtype == MERCURY_F32 ? 4 : 2, row_bytes = a.c * elt;
    const int nchunks = (int)((a.n + kChunk - 1) / kChunk);
    if (row_bytes > 256 || nchunks > kStepMaxChunks) return 1;
    if (fused_mode < 0 && nchunks > kStepBestChunks) return 1;   // (forced on by the test hook it takes all it can)
    if ((reinterpret_cast<uintptr_t>(a.logits) & 15) || (reinterpret_cast<uintptr_t>(a.labels) & 15)) return 1;
    if (walk2_slots(nchunks) < 4) return 1;
    const CdfWs w = carve(ws, a.n, true);
    StepArgs sa{};
    sa.logits = a.logits; sa.labels = a.labels; sa.n = a.n; sa.c = (int)a.c; sa.tile_rows = small_tile_rows(row_bytes);
    sa.ignore_index = a.ignore_index; sa.scores = a.scores; sa.part = w.sums;
    sa.ema_state = a.ema_state; sa.ema_first = a.ema_first; sa.ema_alpha = a.ema_alpha; sa.smooth_alpha = a.smooth_alpha; sa.stats = a.stats;
    sa.nchunks = nchunks; sa.bias = g_guess_bias.load(); sa.sums = w.summaries; sa.packs = reinterpret_cast<Pack2*>(w.subs);
    sa.start_bits = w.start_bits; sa.total = w.total; sa.ctrl = w.ctrl; sa.status = a.status;
    sa.cdf_mode = a.cdf_mode; sa.approx = w.approx;
    if (a.peer_boxes) {
        sa.peers = reinterpret_cast<PeerBox* const*>(a.peer_boxes); sa.nshards = a.n_shards; sa.rank = a.shard_rank;
        sa.n_global = a.n_global; sa.sums_all = a.sums_all; sa.totals_all = a.totals_all; sa.local_u = a.local_uniforms;
    }
    static const int dbg_flags = env_knob("MERCURY_STEP_DBG", 0, 0, 255);
    sa.dbg_flags = dbg_flags;
    if (small_smem_bytes(kStepStages, sa.tile_rows, row_bytes) > kStepSmemBytes) return 1;
    DrawArgs& d = sa.d;
    d.n_weight = a.peer_boxes ? a.n_global : a.n; d.idx_base = a.peer_boxes ? a.index_base : 0;
    d.src = WeightSrc{a.scores, a.stats}; d.n = a.n; d.nchunks = nchunks; d.start_bits = w.start_bits; d.total_ptr = w.total;
    d.u = a.uniforms; d.b = a.b;
    d.data = static_cast<const unsigned char*>(a.data); d.row_bytes = a.row_bytes; d.labels = a.labels; d.index = a.index;
    d.idx = a.idx; d.nprob = a.nprob; d.seed = a.seed; d.out_data = static_cast<unsigned char*>(a.out_data);
    d.out_labels = a.out_labels; d.out_index = a.out_index;
    d.approx = w.approx;
    d.dbg = (dbg_flags & 4) ? &w.ctrl->evt[107] : nullptr;
    const int64_t ntiles = (a.n + sa.tile_rows - 1) / sa.tile_rows;
    sa.ntiles = (int)ntiles;
    // CTAs wanted: one per tile of scores, per chunk (+ the walker, which owns none) or per draw, whichever is largest
    int64_t want = ntiles > a.b ? ntiles : a.b;
    if (want < nchunks + 1) want = nchunks + 1;
    if (a.dtype == MERCURY_F32) {
        if (a.c == 10) return fused_launch_typed<float, 10>(sa, want, st);
        return fused_launch_typed<float, 0>(sa, want, st);
    }
    if (a.c == 10) return fused_launch_typed<__nv_bfloat16, 10>(sa, want, st);
    return fused_launch_typed<__nv_bfloat16, 0>(sa, want, st);
}

// developer aid: the stamps the last fused step left in this workspace (16 phase times + 112 walk steps); synchronises
int fused_step_times(const void* ws, int64_t n, unsigned long long* h_out128) {
    const CdfWs w = carve(const_cast<void*>(ws), n, true);
    cudaError_t e = cudaMemcpy(h_out128, w.ctrl->times, 128 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && getenv("MERCURY_STEP_DUMP_K1")) {   // developer aid: per-CTA scoring times to stderr
        static unsigned int h[kStepMaxCtas];
        e = cudaMemcpy(h, w.ctrl->k1_end, sizeof(h), cudaMemcpyDeviceToHost);
        const unsigned int t0 = (unsigned int)h_out128[0];
        for (int i = 0; i < 444 && e == cudaSuccess; ++i) fprintf(stderr, "%d:%.2f ", i, (double)(h[i] - t0) / 1e3);
        fprintf(stderr, "\n");
    }
    return e == cudaSuccess ? MERCURY_OK : cuda_fail(e, "fused_step_times");
}

double* update_ws_part_sums(void* ws, int64_t n) { return carve(ws, n, true).sums; }

}  // namespace mercury

using namespace mercury;

extern "C" void mercury_test_set_guess_bias(double bias) { g_guess_bias = bias; }
extern "C" void mercury_test_set_fused(int mode) { g_fused_mode = mode; }
extern "C" int mercury_debug_step_times(const void* ws, int64_t n, unsigned long long* h_out128) {
    MB_CHECK_ARG(ws && n >= 1 && h_out128, "mercury_debug_step_times: bad arguments");
    return fused_step_times(ws, n, h_out128);
}

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
                             static_cast<cudaStream_t>(stream), false);
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

namespace {
int cdf_run(const char* who, const float* weights, int64_t n, int mode, float* cdf, float* total, int32_t* status, void* ws,
            size_t ws_bytes, int slots, void* stream) {
    MB_CHECK_ARG(n >= 1, "%s: empty distribution", who);
    MB_CHECK_ARG(n <= ((int64_t)1 << 24), "number of categories cannot exceed 2^24");
    MB_CHECK_ARG(weights && cdf && total && status, "%s: null pointer", who);
    MB_CHECK_ARG(mode == MERCURY_CDF_EXACT_SERIAL_F32 || mode == MERCURY_CDF_F64, "%s: unknown mode %d", who, mode);
    const CdfWs w = carve(ws, n, false, slots);
    int rc = check_ws(ws, ws_bytes, w.bytes, who);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    chunk_sums_kernel<<<nchunks, 256, 0, st>>>(weights, n, w.sums, status, 1);
    MB_CHECK_LAUNCH("chunk_sums_kernel");
    prefix_finalize_kernel<<<1, 1024, 0, st>>>(w.sums, nchunks, w.approx, w.slotid, g_guess_bias.load(), slots);
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
    if (e != cudaSuccess) return cuda_fail(e, "cdf: total copy");
    return MERCURY_OK;
}
}  // namespace

extern "C" int mercury_cdf(const float* weights, int64_t n, int mode, float* cdf, float* total, int32_t* status, void* ws,
                           size_t ws_bytes, void* stream) {
    return cdf_run("mercury_cdf", weights, n, mode, cdf, total, status, ws, ws_bytes, kSubSlots, stream);
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
// Sharded pools, "totals only" exchange (mode B).
namespace mercury {
namespace {
// deterministic sum of the chunk sums (single CTA, fixed order)
__global__ void __launch_bounds__(1024) reduce_sums_kernel(const double* __restrict__ sums, int nchunks, double* __restrict__ out) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < nchunks; i += 1024) acc += sums[i];
    const double t = block_sum<1024>(acc, red);
    if (threadIdx.x == 0) *out = t;
}

// first level of the two-level draw: which shard a uniform falls in (serial fp32 CDF over the shard
// totals, searched as ATen does), and where inside that shard's own CDF
__global__ void __launch_bounds__(256) shard_select_kernel(const float* __restrict__ shard_totals, int nshards, int rank,
                                                           const double* __restrict__ u, int64_t b, double* __restrict__ local_u) {
    __shared__ float cdf[kMaxShards];
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int g = 0; g < nshards; ++g) { s = __fadd_rn(s, shard_totals[g]); cdf[g] = s; }
        for (int g = 0; g < nshards; ++g) cdf[g] = __fdiv_rn(cdf[g], s);
    }
    __syncthreads();
    for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < b; j += (int64_t)gridDim.x * 256) {
        const double uj = u[j];
        int g = 0;
        while (g < nshards - 1 && (double)cdf[g] < uj) ++g;   // first shard with cdf >= u (last one catches the rest)
        double out = -1.0;
        if (g == rank) {
            const double lo = g > 0 ? (double)cdf[g - 1] : 0.0, hi = (double)cdf[g];
            out = hi > lo ? (uj - lo) / (hi - lo) : 0.0;
            if (out < 0.0) out = 0.0;
            if (out >= 1.0) out = 0.99999999999999989;  // largest double below 1
        }
        local_u[j] = out;
    }
}

// (PeerBox and its helpers are defined near the top of this file, before FusedStats)
__global__ void __launch_bounds__(1024) sum_exchange_kernel(const double* __restrict__ sums, int nchunks, PeerBox* const* __restrict__ peers,
                                                            int nshards, int rank, double* __restrict__ sums_all,
                                                            int32_t* __restrict__ status) {
    __shared__ double red[32];
    __shared__ unsigned int s_seq;
    __shared__ double s_sum;
    pdl_wait();
    pdl_launch_dependents();
    double acc = 0.0;
    for (int i = threadIdx.x; i < nchunks; i += 1024) acc += sums[i];
    const double t = block_sum<1024>(acc, red);
    PeerBox* mine = peers[rank];
    if (threadIdx.x == 0) {
        s_seq = ++mine->seq_sum;
        s_sum = t;
    }
    __syncthreads();
    const unsigned int seq = s_seq;
    const int par = seq & 1u, g = threadIdx.x;
    if (g < nshards) {
        const unsigned long long bits = (unsigned long long)__double_as_longlong(s_sum);
        peer_store(&peers[g]->sum_hi[par][rank], (bits >> 32 << 32) | seq);
        peer_store(&peers[g]->sum_lo[par][rank], (bits << 32) | seq);
        unsigned long long hi, lo;
        const bool ok = peer_wait(&mine->sum_hi[par][g], seq, hi, mine) & peer_wait(&mine->sum_lo[par][g], seq, lo, mine);
        if (!ok) atomicOr(status, MERCURY_ST_PEER_TIMEOUT);
        sums_all[g] = __longlong_as_double((long long)((hi >> 32 << 32) | (lo >> 32)));
    }
}

__global__ void __launch_bounds__(256) select_exchange_kernel(const float* __restrict__ my_total, PeerBox* const* __restrict__ peers,
                                                              int nshards, int rank, const double* __restrict__ u, int64_t b,
                                                              float* __restrict__ totals_all, double* __restrict__ local_u,
                                                              int32_t* __restrict__ status) {
    __shared__ float tot[kMaxShards];
    __shared__ float cdf[kMaxShards];
    __shared__ unsigned int s_seq;
    pdl_wait();
    pdl_launch_dependents();
    PeerBox* mine = peers[rank];
    if (threadIdx.x == 0) s_seq = ++mine->seq_total;
    __syncthreads();
    const unsigned int seq = s_seq;
    const int par = seq & 1u, g = threadIdx.x;
    if (g < nshards) {
        peer_store(&peers[g]->total[par][rank], ((unsigned long long)__float_as_uint(*my_total) << 32) | seq);
        unsigned long long w;
        if (!peer_wait(&mine->total[par][g], seq, w, mine)) atomicOr(status, MERCURY_ST_PEER_TIMEOUT);
        tot[g] = __uint_as_float((unsigned int)(w >> 32));
        totals_all[g] = tot[g];
    }
    __syncthreads();
    const bool bad = (*reinterpret_cast<volatile int32_t*>(status) & MERCURY_ST_PEER_TIMEOUT) != 0;
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int k = 0; k < nshards; ++k) { s = __fadd_rn(s, tot[k]); cdf[k] = s; }
        for (int k = 0; k < nshards; ++k) cdf[k] = __fdiv_rn(cdf[k], s);
    }
    __syncthreads();
    for (int64_t j = threadIdx.x; j < b; j += 256) {
        const double uj = u[j];
        int k = 0;
        while (k < nshards - 1 && (double)cdf[k] < uj) ++k;
        double out = -1.0;
        if (k == rank) {
            const double lo = k > 0 ? (double)cdf[k - 1] : 0.0, hi = (double)cdf[k];
            out = hi > lo ? (uj - lo) / (hi - lo) : 0.0;
            if (out < 0.0) out = 0.0;
            if (out >= 1.0) out = 0.99999999999999989;
        }
        local_u[j] = bad ? peer_poison() : out;
    }
}
}  // namespace
}  // namespace mercury

namespace mercury {
namespace {
// wait_ns > 0: set the bound; reset: back to the state mercury_peer_create leaves (words, step counters and `dead` cleared)
__global__ void peer_config_kernel(PeerBox* box, unsigned long long wait_ns, int reset) {
    if (reset) {
        unsigned long long* w = reinterpret_cast<unsigned long long*>(box);
        for (int i = threadIdx.x; i < (int)(offsetof(PeerBox, wait_ns) / 8); i += blockDim.x) w[i] = 0ull;
    }
    if (threadIdx.x == 0 && wait_ns) box->wait_ns = wait_ns;
}
static_assert(offsetof(PeerBox, wait_ns) % 8 == 0, "the words in front of wait_ns are cleared 8 bytes at a time");
}  // namespace
}  // namespace mercury

extern "C" int mercury_peer_set_wait(void* box, int64_t wait_ms, void* stream) {
    MB_CHECK_ARG(box && wait_ms > 0 && wait_ms <= 3600000, "mercury_peer_set_wait: box must be set and 0 < wait_ms <= 3 600 000");
    mercury::peer_config_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(static_cast<mercury::PeerBox*>(box),
                                                                    (unsigned long long)wait_ms * 1000000ull, 0);
    MB_CHECK_LAUNCH("peer_config_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_peer_reset(void* box, void* stream) {
    MB_CHECK_ARG(box, "mercury_peer_reset: null pointer");
    mercury::peer_config_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(static_cast<mercury::PeerBox*>(box), 0ull, 1);
    MB_CHECK_LAUNCH("peer_config_kernel");
    return MERCURY_OK;
}

namespace mercury {
// the whole sharded step behind mercury_update_samples (args.peer_boxes set): one PDL chain
int sharded_step_launch(const mercury_update_args& a, const float* scores, const double* part, int per, void* ws, cudaStream_t st) {
    // Four kernels, as on one GPU: score (done by the caller) -> summarize (pushes this shard's sum of scores, waits for all
    // G) -> spine (publishes the shard's chain total) -> draw (waits for the G totals, picks the shard of each uniform,
    // resolves the ones that are its own).  No kernel exists only to communicate.
    const CdfWs w = carve(ws, a.n, true);
    const int nchunks = (int)((a.n + kChunk - 1) / kChunk);
    if (per == 0) {
        chunk_sums_kernel<<<nchunks, 256, 0, st>>>(scores, a.n, w.sums, nullptr, 0);
        MB_CHECK_LAUNCH("chunk_sums_kernel");
        part = w.sums;
        per = kChunk;
    }
    PeerBox* const* peers = reinterpret_cast<PeerBox* const*>(a.peer_boxes);
    WeightSrc src{scores, a.stats};
    FusedStats fs{};
    fs.part = part; fs.per = per; fs.npart = (int)((a.n + per - 1) / per);
    fs.ema_in = a.ema_state; fs.ema_first = a.ema_first; fs.ema_alpha = a.ema_alpha; fs.smooth_alpha = a.smooth_alpha;
    fs.stats = a.stats;
    fs.n_global = a.n_global;
    fs.peers = peers; fs.nshards = a.n_shards; fs.rank = a.shard_rank; fs.gsums_out = a.sums_all;
    fs.gsum = w.ctrl->gsum_chain;
    int rc = run_exact_boundaries(src, a.n, w, a.status, 1, st, &fs, a.ema_state);
    if (rc) return rc;
    if (a.b > 0) {
        DrawArgs d{};
        d.n_weight = a.n_global; d.idx_base = a.index_base;
        d.src = src; d.n = a.n; d.nchunks = nchunks; d.start_bits = w.start_bits; d.total_ptr = w.total;
        d.peers = peers; d.nshards = a.n_shards; d.rank = a.shard_rank; d.totals_all = a.totals_all; d.local_u = a.local_uniforms;
        d.status = a.status;
        d.u = a.uniforms; d.b = a.b;
        d.data = static_cast<const unsigned char*>(a.data); d.row_bytes = a.row_bytes; d.labels = a.labels; d.index = a.index;
        d.idx = a.idx; d.nprob = a.nprob; d.seed = a.seed; d.out_data = static_cast<unsigned char*>(a.out_data);
        d.out_labels = a.out_labels; d.out_index = a.out_index;
        cudaError_t e = launch_chain(true, draw_kernel<true>, dim3((unsigned)a.b), dim3(kDrawNT), 0, st, d);
        if (e != cudaSuccess) return cuda_fail(e, "draw_kernel launch");
    }
    return MERCURY_OK;
}
}  // namespace mercury

extern "C" int mercury_sum_scores_exchange(const float* scores, int64_t n, void* const* peer_boxes, int nshards, int rank,
                                           double* sums_all, int32_t* status, void* ws, size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1 && scores && sums_all && peer_boxes && status, "mercury_sum_scores_exchange: bad arguments");
    MB_CHECK_ARG(nshards >= 1 && nshards <= kMaxShards && rank >= 0 && rank < nshards, "mercury_sum_scores_exchange: 1..%d shards", kMaxShards);
    const CdfWs w = carve(ws, n, true);
    int rc = check_ws(ws, ws_bytes, w.bytes, "mercury_sum_scores_exchange");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    chunk_sums_kernel<<<nchunks, 256, 0, st>>>(scores, n, w.sums, nullptr, 0);
    MB_CHECK_LAUNCH("chunk_sums_kernel");
    sum_exchange_kernel<<<1, 1024, 0, st>>>(w.sums, nchunks, reinterpret_cast<PeerBox* const*>(peer_boxes), nshards, rank, sums_all, status);
    MB_CHECK_LAUNCH("sum_exchange_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_shard_select_exchange(const float* my_total, void* const* peer_boxes, int nshards, int rank,
                                             const double* uniforms, int64_t b, float* totals_all, double* local_uniforms,
                                             int32_t* status, void* stream) {
    MB_CHECK_ARG(nshards >= 1 && nshards <= kMaxShards, "mercury_shard_select_exchange: 1..%d shards", kMaxShards);
    MB_CHECK_ARG(rank >= 0 && rank < nshards && b >= 0, "mercury_shard_select_exchange: bad rank or draw count");
    MB_CHECK_ARG(my_total && peer_boxes && totals_all && status && (b == 0 || (uniforms && local_uniforms)),
                 "mercury_shard_select_exchange: null pointer");
    select_exchange_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(my_total, reinterpret_cast<PeerBox* const*>(peer_boxes), nshards,
                                                                             rank, uniforms, b, totals_all, local_uniforms, status);
    MB_CHECK_LAUNCH("select_exchange_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_sum_scores(const float* scores, int64_t n, double* out_sum, void* ws, size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(n >= 1 && scores && out_sum, "mercury_sum_scores: bad arguments");
    const CdfWs w = carve(ws, n, true);
    int rc = check_ws(ws, ws_bytes, w.bytes, "mercury_sum_scores");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nchunks = (int)((n + kChunk - 1) / kChunk);
    chunk_sums_kernel<<<nchunks, 256, 0, st>>>(scores, n, w.sums, nullptr, 0);
    MB_CHECK_LAUNCH("chunk_sums_kernel");
    reduce_sums_kernel<<<1, 1024, 0, st>>>(w.sums, nchunks, out_sum);
    MB_CHECK_LAUNCH("reduce_sums_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_shard_select(const float* shard_totals, int nshards, int rank, const double* uniforms, int64_t b,
                                    double* local_uniforms, void* stream) {
    MB_CHECK_ARG(nshards >= 1 && nshards <= kMaxShards, "mercury_shard_select: 1..%d shards", kMaxShards);
    MB_CHECK_ARG(rank >= 0 && rank < nshards && b >= 0, "mercury_shard_select: bad rank or draw count");
    if (b == 0) return MERCURY_OK;
    MB_CHECK_ARG(shard_totals && uniforms && local_uniforms, "mercury_shard_select: null pointer");
    shard_select_kernel<<<(unsigned)((b + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(shard_totals, nshards, rank, uniforms, b,
                                                                                                    local_uniforms);
    MB_CHECK_LAUNCH("shard_select_kernel");
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
// every GPU that loads the library carries these 208 KiB until the process exits.
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
constexpr size_t kStaticScratchBytes = 208 * 1024;
constexpr int kStaticSlots = 24;  // packs for chunks that change binade; beyond that the exact fallback path is taken
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
    if (carve(nullptr, n, false, kStaticSlots).bytes > kStaticScratchBytes) {  // cannot happen for n <= kStaticMaxN; a guard
        set_last_error("mercury_cdf_static: internal scratch too small for n = %lld", (long long)n);
        return MERCURY_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = static_scratch_for_current_device(&scratch);
    if (rc) return rc;
    return cdf_run("mercury_cdf_static", weights, n, mode, cdf, total, status, scratch, kStaticScratchBytes, kStaticSlots, stream);
}
