// The whole step as ONE persistent cooperative kernel (included by cdf.cu, inside its anonymous namespace).
//
// Replaces the arithmetic of Trainer.update_samples (reference pytorch_collab.py:102-117) for pools of small rows
// (<= 256 B of logits per candidate: the reference's 10-class head).  The four-kernel chain (score -> summarize ->
// spine -> draw) spends most of its time between kernels and in a one-CTA serial walk; here the same work runs as
// phases of one grid that stays resident:
//   A  K1 (score_small_body): scores + the fp64 sum of every tile of scores;
//      -- grid-wide barrier (cooperative groups) --
//   B  every CTA: pool statistics (mean / EMA / shift / total, bit-identical in all CTAs), then for each of its chunks
//      the 2-state map of the chunk in the guessed phase; a chunk in which the running sum may change binade also gets
//      a Pack2: one map per 32-candidate sub-chunk for the guessed phase and the next, and the raw candidates of the
//      sub-chunks in which the change should fall;
//      -- the last CTA is the walker: it owns no chunks and polls the summaries (tagged with the launch's start time) --
//   C  the walker: loads all summaries + packs into shared memory (one round trip), walks the chain (chunk level as in
//      spine_kernel; inside a pack the 64 sub-chunk increments are scanned by the warp, the first sub-chunk whose end
//      leaves the phase is summed with the reference's own loop, 32 dependent fp32 adds), writes the exact boundary
//      values, finds the chunk of every uniform and publishes (chunk, boundary value) per draw;
//   D  every CTA: polls the word of its draws, rebuilds that chunk's prefix (draw_resolve) and gathers the row.
// Results are bit-identical to the four-kernel chain (same statistics, same p, same chain); a contradicted guess falls
// back to the same general code (serial_chunk).
#pragma once
// (cdf.cu includes <cooperative_groups.h> at global scope)
namespace cg = cooperative_groups;

constexpr int kStepNT = 256;                     // worker threads: warps 0..7 (K1's consumers); warp 8 is K1's producer
constexpr int kStepThreads = kStepNT + kProducerThreads;
constexpr int kStepStages = 3;
constexpr int kStepMaxChunks = 1024;             // pools of up to 2 Mi candidates (shared-memory layout)
constexpr int kStepBestChunks = 768;             // ... taken by default up to 1.5 Mi (measured: the kernel chain is faster beyond)
constexpr int kStepMaxDraws = 1024;
constexpr int kStepMaxCtas = 1024;
constexpr int kStepMaxPacks = 64;                // flagged chunks that get a pack (beyond: the general code)
constexpr int kStepMaxEvents = 192;
constexpr size_t kStepSmemBytes = 70 * 1024;
constexpr int kSub2 = 32, kSubs2 = kChunk / kSub2;
constexpr int kRaw2 = 12;
constexpr int kHeadMax = 1024;                   // candidates at the head of the pool summed serially during phase B
constexpr uint32_t kBadInc = kSatCap;            // "no plain increment known for this phase": forces a stop in the scan
static_assert(kSubs2 == 64, "two sub-chunks per lane of the walking warp");

struct __align__(128) Pack2 {
    uint32_t e[kSubs2];            // guessed phase | kFlag (the change may fall here) | kTie0 | kTie1; kAnyPhase: all zero
    uint2 q0[kSubs2];              // the sub-chunk's map in its guessed phase
    uint2 q1[kSubs2];              // ... and in the phase above
    uint8_t rawsub[16];            // sub-chunks whose candidates are in raw[] (0xff: unused)
    uint32_t raw[kRaw2][kSub2];    // 16-byte aligned
    // inclusive prefix sums of the sub-chunks' plain increments while the running sum is in phase e_lo (psA) or e_lo + 1
    // (psB); a sub-chunk without a plain increment for that phase counts kBadInc, which pushes every later entry beyond
    // 2^24 (so it reads as "leaves the phase here")
    uint32_t psA[kSubs2], psB[kSubs2];
    uint32_t e_lo;                 // the chunk's low guess
    uint32_t pad_[31];
};
static_assert(sizeof(Pack2) == 3584 && offsetof(Pack2, raw) % 16 == 0 && offsetof(Pack2, q0) % 16 == 0 && offsetof(Pack2, psA) % 16 == 0,
              "Pack2 layout");

struct StepCtrl {
    unsigned int nflag;            // pack slots handed out
    unsigned int arrive_unused;
    unsigned int packs_done;       // packs written (the walker waits for min(nflag, kStepMaxPacks))
    unsigned int pad_[27];
    unsigned long long epoch;      // this launch's tag (%globaltimer at its start): a summary is valid once it carries it

    unsigned long long times[16];  // developer aid: %globaltimer at the phase changes (CTA 0 and the walker)
    unsigned long long evt[112];   // developer aid: clock64 after each of the walk's first steps (kind in the low 4 bits)
    unsigned long long dinfo[kStepMaxDraws];  // per draw: 1 << 63 | chunk << 32 | boundary bits
    unsigned long long dinfo2[kStepMaxDraws]; // per draw: 1 << 63 | bits of the chain's total (each word is valid on its own:
                                              // no fence between the walker's results and these words is needed)
    double cta_part[kStepMaxCtas];            // per CTA: the fp64 sum of the scores it computed (its tiles are contiguous)
    double dapprox[kStepMaxDraws];            // fp64 mode, per draw: the running total of p in front of its chunk ...
    double dtotal[kStepMaxDraws];             // ... and the total (written before dinfo[j], a fence in between)
    unsigned int k1_end[kStepMaxCtas];        // developer aid (MERCURY_STEP_DBG & 4): ns from the start until each CTA had scored
    // sharded pools: the whole pool's sum of scores, received from the peers by CTA 0 and handed to the other CTAs in two
    // self-validating words (1 << 63 | half of the double's bits) -- 443 CTAs polling the peers' words themselves was a storm of
    // system-scope loads on one cache line; and, per draw, 1 << 63 | bits of the uniform rescaled into this shard
    unsigned long long gsum[2];
    unsigned long long gsum_chain[2];   // the same hand-over in the kernel chain's summarize_kernel (half of the bits << 32 | exchange tag)
    unsigned long long dinfo3[kStepMaxDraws];
};
constexpr unsigned long long kStepSpinNs = 2000000000ull;  // no wait may last this long: report instead of hanging the GPU

struct StepArgs {
    // K1
    const void* logits; const int64_t* labels; int64_t n; int c; int tile_rows; int ntiles; int chunks_q, chunks_r, parts_per_chunk; int64_t ignore_index; float* scores;
    double* part;                  // [ntiles] fp64 sums of the tiles' scores
    // statistics
    float* ema_state; int ema_first; double ema_alpha, smooth_alpha; float* stats;
    // chain
    int nchunks; double bias; Summary* sums; Pack2* packs; uint32_t* start_bits; float* total; StepCtrl* ctrl;
    int32_t* status;
    int dbg_flags;                 // developer aid (MERCURY_STEP_DBG): 1 = stamp the walk's steps
    // sharded pool, exchanges inside the kernel (peers != nullptr; see mercury_peer_create): this launch covers one shard
    PeerBox* const* peers; int nshards, rank; int64_t n_global; double* sums_all; float* totals_all; double* local_u;
    int cdf_mode;                  // MERCURY_CDF_EXACT_SERIAL_F32 or MERCURY_CDF_F64 (no walk: fp64 chunk sums and their scan)
    double* approx;                // fp64 mode: [nchunks + 1] running total of p in front of every chunk
    // draws
    DrawArgs d;
};

__device__ __forceinline__ void wsync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint2 lds_u32x2(uint32_t saddr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint4 lds_u32x4(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}

// the increment of a sub-chunk while the running sum is in phase E, or kBadInc if no plain one is known (branch-free)
__device__ __forceinline__ uint32_t sub_inc(uint32_t e, uint32_t d00, uint32_t d01, uint32_t d10, uint32_t d11, uint32_t E) {
    const uint32_t g = e & kPhaseMask;
    const uint32_t v0 = d00 == d01 ? d00 : kBadInc, v1 = d10 == d11 ? d10 : kBadInc;
    const uint32_t r = g == E ? v0 : (g + 1u == E ? v1 : kBadInc);
    return g == kAnyPhase ? 0u : r;
}

// Sum of the plain increments of 8 addends while the running sum is in phase E, and whether any of them is an exact tie
// (then the sum is meaningless and the caller composes the 2-state maps instead).  p / ulp is p times a power of two,
// exact in fp32, and away from ties the increment is that number rounded to nearest: 7 instructions per addend.
__device__ __forceinline__ void plain_incs8(const uint32_t (&pb)[8], uint32_t E, uint32_t& sum, bool& tie) {
    sum = 0u;
    tie = false;
    if (E < kFastPhaseMin) {   // (running sums below 2^-104: the scale is not a normal float; the integer version decides)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const Pair q = addend_pair_fast(pb[k], E);
            tie |= q.d0 != q.d1;
            sum += q.d0;
        }
        return;
    }
    const float scale = b2f((276u - E) << 23);             // 2^(149 - E)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float t = fminf(b2f(pb[k]) * scale, 33554432.f);   // capped at 2^25 (inf and NaN land there too)
        const float r = rintf(t);
        sum += (uint32_t)r;                                 // (8 * 2^25 fits)
        tie |= fabsf(r - t) == 0.5f;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// phase B: one chunk -> Summary (+ Pack2)
struct Sum2Smem {
    Pair wt[kStepNT / 32];
    float wsum[kStepNT / 32];
    int wnz[kStepNT / 32];
    unsigned int wflag[kStepNT / 32];
    double dred[2][kStepNT / 32];
    double own[8];
    int slot;
    uint32_t sub_e[kSubs2]; uint2 sub_q0[kSubs2], sub_q1[kSubs2];
};

// epoch: this launch's tag.  The summary carries it in both of its 16-byte halves (each written by one vector store), which
// is how the walker knows that a summary it reads is this launch's: no fence, no counter.
__device__ __forceinline__ void summarize2_chunk(const StepArgs& a, int c, const float (&raw)[8], double pre, double own, float shift,
                                                 float total, Sum2Smem& sm, unsigned long long epoch, bool first) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t start = (int64_t)c * kChunk;
    const int len = (int)min((int64_t)kChunk, a.n - start);
    const double inv_total = 1.0 / (double)total;   // (the approximate running totals only feed the guesses)
    const double a_lo = (pre + (double)start * (double)shift) * inv_total;
    const double a_hi = a_lo + (own + (double)len * (double)shift) * inv_total;
    uint32_t pb[8];
    {
        const int base = threadIdx.x * 8;
#pragma unroll
        for (int k = 0; k < 8; ++k) pb[k] = base + k < len ? f2b(__fdiv_rn(__fadd_rn(raw[k], shift), total)) : 0u;
    }
#define MB_BSTAMP(i) do { if ((a.dbg_flags & 4) && c == 0 && threadIdx.x == 64) a.ctrl->evt[100 + (i)] = global_ns(); } while (0)
    MB_BSTAMP(0);
    uint32_t e;
    const bool flagged = guess_span(a_lo, a_hi, a.bias, &e, chunk_margin(a.nchunks));  // block-uniform
    // the usual flagged chunk meets ONE change of binade: its sub-chunks are then all summarised for the chunk's guessed
    // phase and the next, whatever their own position (the head of the pool, with many changes, guesses per sub-chunk)
    const bool two_phase = flagged && phase_of(f2b((float)(a_hi * a.bias * (1.0 + chunk_margin(a.nchunks))))) == e + 1u;
    // a flagged chunk claims a pack slot; the atomic's round trip hides behind the arithmetic below
    unsigned int slot_u = 0xffffffffu;
    if (flagged && threadIdx.x == 0) slot_u = atomicAdd(&a.ctrl->nflag, 1u);
    Pair th0, th1;   // this thread's 8 addends as a map, in phase e (and e + 1: two-phase chunks only)
    int bad = 0;
    bool nz = false;
    float tsum = 0.f;
    {
        // without exact ties every addend is a plain increment and the thread's map is their sum (the usual case);
        // only a thread that meets a tie composes the 2-state maps
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float x = b2f(pb[k]);
            if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
            else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
            nz |= pb[k] != 0u;
            tsum += x;
        }
        uint32_t plain;
        bool tie;
        plain_incs8(pb, e, plain, tie);
        th0.d0 = th0.d1 = sat(plain);
        if (tie || bad) {
            th0 = identity_pair();
#pragma unroll
            for (int k = 0; k < 8; ++k) th0 = compose(th0, addend_pair_fast(pb[k], e));
        }
        th1 = th0;
        if (two_phase) {
            plain_incs8(pb, e + 1u, plain, tie);
            th1.d0 = th1.d1 = sat(plain);
            if (tie || bad) {
                th1 = identity_pair();
#pragma unroll
                for (int k = 0; k < 8; ++k) th1 = compose(th1, addend_pair_fast(pb[k], e + 1u));
            }
        }
    }
    if (bad) atomicOr(a.status, bad);
    const unsigned nzmask = __ballot_sync(0xffffffffu, nz);
    Pair a0 = th0;
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
    float inc = tsum;  // inclusive scan of the thread sums (approximate: only feeds the sub-chunk guesses)
    if (flagged) {
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const float o = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += o;
        }
    }
    MB_BSTAMP(1);
    if (!first) wsync();  // (protects sm against the previous chunk's readers)
    if (lane == 0) { sm.wt[warp] = a0; sm.wnz[warp] = nzmask != 0u; }
    if (lane == 31) sm.wsum[warp] = inc;
    if (threadIdx.x == 0) sm.slot = flagged ? (slot_u < (unsigned)kStepMaxPacks ? (int)slot_u : -2) : -1;
    wsync();
    const int slot = sm.slot;
    if (warp == 0) {   // the chunk's map: the 8 warp maps in order (lanes 0..7), one integer reduction unless there are ties
        Pair t0 = lane < kStepNT / 32 ? sm.wt[lane] : identity_pair();
        const unsigned any = __ballot_sync(0xffffffffu, lane < kStepNT / 32 && sm.wnz[lane] != 0);
        if (__ballot_sync(0xffffffffu, t0.d0 != t0.d1) == 0u) {
            t0.d0 = t0.d1 = sat(__reduce_add_sync(0xffffffffu, t0.d0));
        } else {
#pragma unroll
            for (int off = 1; off < kStepNT / 32; off <<= 1) {
                Pair o0;
                o0.d0 = __shfl_down_sync(0xffffffffu, t0.d0, off); o0.d1 = __shfl_down_sync(0xffffffffu, t0.d1, off);
                if (lane + off < kStepNT / 32) t0 = compose(t0, o0);
            }
        }
        if (lane == 0) {
            const uint32_t se = (any ? e : kAnyPhase) | (slot >= 0 ? kFlag : 0u) | (flagged && slot < 0 ? kNoPack : 0u) | (t0.d0 != t0.d1 ? kTie0 : 0u);
            uint4* dst = reinterpret_cast<uint4*>(a.sums + c);
            // {e, q0.d0, q0.d1, tag lo} {tag hi, slot, 0, 0}: Summary's q[1] and pad[1..] are not used on this path
            dst[0] = make_uint4(se, t0.d0, t0.d1, (uint32_t)epoch);
            dst[1] = make_uint4((uint32_t)(epoch >> 32), (uint32_t)slot, 0u, 0u);
        }
    }
    if (slot < 0) return;
    MB_BSTAMP(2);
    // ---- per sub-chunk: 4 consecutive threads = 32 candidates
    float pre_f = 0.f;
    for (int i = 0; i < warp; ++i) pre_f += sm.wsum[i];
    const float t_lo = pre_f + inc - tsum, t_hi = pre_f + inc;
    const int lead = lane & ~3;
    const float s_lo = __shfl_sync(0xffffffffu, t_lo, lead), s_hi = __shfl_sync(0xffffffffu, t_hi, lead + 3);
    const bool sub_nz = ((nzmask >> lead) & 0xfu) != 0u;
    uint32_t es;
    const bool sub_flagged = guess_span(a_lo + (double)s_lo, a_lo + (double)s_hi, a.bias, &es, sub_margin(a.nchunks)) && sub_nz;
    if (two_phase) es = e;
    Pair m0 = th0, m1 = th1;
    if (!two_phase) {
        uint32_t p0, p1;
        bool tie0, tie1;
        plain_incs8(pb, es, p0, tie0);
        plain_incs8(pb, es + 1u, p1, tie1);
        const bool tie = tie0 | tie1 | (bad != 0);
        m0.d0 = m0.d1 = sat(p0); m1.d0 = m1.d1 = sat(p1);
        if (tie) {
            m0 = identity_pair(); m1 = identity_pair();
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                m0 = compose(m0, addend_pair_fast(pb[k], es));
                m1 = compose(m1, addend_pair_fast(pb[k], es + 1u));
            }
        }
    }
    if (__ballot_sync(0xffffffffu, (m0.d0 != m0.d1) | (m1.d0 != m1.d1)) == 0u) {   // no ties in this warp: plain sums
        uint32_t v0 = m0.d0, v1 = m1.d0;
#pragma unroll
        for (int off = 1; off < 4; off <<= 1) {
            v0 += __shfl_down_sync(0xffffffffu, v0, off, 4);
            v1 += __shfl_down_sync(0xffffffffu, v1, off, 4);
        }
        m0.d0 = m0.d1 = sat(v0); m1.d0 = m1.d1 = sat(v1);   // (meaningful in the leader lanes)
    } else {
#pragma unroll
        for (int off = 1; off < 4; off <<= 1) {
            Pair o0, o1;
            o0.d0 = __shfl_down_sync(0xffffffffu, m0.d0, off, 4); o0.d1 = __shfl_down_sync(0xffffffffu, m0.d1, off, 4);
            o1.d0 = __shfl_down_sync(0xffffffffu, m1.d0, off, 4); o1.d1 = __shfl_down_sync(0xffffffffu, m1.d1, off, 4);
            if ((lane & 3) + off < 4) { m0 = compose(m0, o0); m1 = compose(m1, o1); }
        }
    }
    MB_BSTAMP(3);
    Pack2* pk = a.packs + slot;
    const int my_sub = threadIdx.x >> 2;
    if ((lane & 3) == 0) {
        // (the word is completed below, once the raw slots are known: bits 10..13 = raw slot + 1)
        pk->q0[my_sub] = make_uint2(m0.d0, m0.d1);
        pk->q1[my_sub] = make_uint2(m1.d0, m1.d1);
        sm.sub_e[my_sub] = sub_nz ? es : kAnyPhase;
        sm.sub_q0[my_sub] = make_uint2(m0.d0, m0.d1);
        sm.sub_q1[my_sub] = make_uint2(m1.d0, m1.d1);
    }
    // which sub-chunks ship their raw candidates: the first kRaw2 flagged ones
    const unsigned lead_flags = __ballot_sync(0xffffffffu, (lane & 3) == 0 && sub_flagged);  // bits 0,4,...,28
    if (lane == 0) {
        unsigned int eight = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) eight |= ((lead_flags >> (4 * i)) & 1u) << i;
        sm.wflag[warp] = eight;
    }
    wsync();
    unsigned long long fmask = 0ull;
#pragma unroll
    for (int i = 0; i < kStepNT / 32; ++i) fmask |= (unsigned long long)sm.wflag[i] << (8 * i);
    const int my_rank = __popcll(fmask & ((1ull << my_sub) - 1ull));
    if (((fmask >> my_sub) & 1ull) && my_rank < kRaw2) {
        uint4* dst = reinterpret_cast<uint4*>(&pk->raw[my_rank][(threadIdx.x & 3) * 8]);
        dst[0] = make_uint4(pb[0], pb[1], pb[2], pb[3]);
        dst[1] = make_uint4(pb[4], pb[5], pb[6], pb[7]);
        if ((lane & 3) == 0) pk->rawsub[my_rank] = (uint8_t)my_sub;
    }
    if ((lane & 3) == 0) {
        const uint32_t raw_slot = (((fmask >> my_sub) & 1ull) && my_rank < kRaw2) ? (uint32_t)(my_rank + 1) : 0u;
        pk->e[my_sub] = (sub_nz ? es : kAnyPhase) | (sub_flagged ? kFlag : 0u) | (m0.d0 != m0.d1 ? kTie0 : 0u) | (m1.d0 != m1.d1 ? kTie1 : 0u) |
                        (raw_slot << 10);
    }
    if (threadIdx.x < 16 && (int)threadIdx.x >= min(__popcll(fmask), kRaw2)) pk->rawsub[threadIdx.x] = 0xff;
    if (warp == 0) {   // prefix sums of the 64 plain increments for the chunk's guessed phase and the next (2 per lane)
        const uint32_t e0 = sm.sub_e[2 * lane], e1 = sm.sub_e[2 * lane + 1];
        const uint2 a0q = sm.sub_q0[2 * lane], a1q = sm.sub_q0[2 * lane + 1], b0q = sm.sub_q1[2 * lane], b1q = sm.sub_q1[2 * lane + 1];
        const uint32_t A0 = sub_inc(e0, a0q.x, a0q.y, b0q.x, b0q.y, e), A1 = sub_inc(e1, a1q.x, a1q.y, b1q.x, b1q.y, e);
        const uint32_t B0 = sub_inc(e0, a0q.x, a0q.y, b0q.x, b0q.y, e + 1u), B1 = sub_inc(e1, a1q.x, a1q.y, b1q.x, b1q.y, e + 1u);
        uint32_t va = A0 + A1, vb = B0 + B1;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t oa = __shfl_up_sync(0xffffffffu, va, off), ob = __shfl_up_sync(0xffffffffu, vb, off);
            if (lane >= off) { va += oa; vb += ob; }
        }
        reinterpret_cast<uint2*>(pk->psA)[lane] = make_uint2(va - A1, va);
        reinterpret_cast<uint2*>(pk->psB)[lane] = make_uint2(vb - B1, vb);
        if (lane == 0) pk->e_lo = e;
    }
    MB_BSTAMP(4);
    // the pack is complete: tell the walker
    wsync();
    if (threadIdx.x == 0) { __threadfence(); atomicAdd(&a.ctrl->packs_done, 1u); }   // (release: cumulative over the barrier)
}

// fp64 mode, phase B: the chunk's fp64 sum of p, as one 16-byte record {sum, launch tag} where its Summary would be
__device__ __forceinline__ void sum_f64_chunk(const StepArgs& a, int c, const float (&raw)[8], float shift, float total, Sum2Smem& sm,
                                              unsigned long long epoch, bool first) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t start = (int64_t)c * kChunk;
    const int len = (int)min((int64_t)kChunk, a.n - start);
    const int base = threadIdx.x * 8;
    double acc = 0.0;
    int bad = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float x = base + k < len ? __fdiv_rn(__fadd_rn(raw[k], shift), total) : 0.f;
        if (!(x >= 0.f)) bad |= MERCURY_ST_NEGATIVE;
        else if (isinf(x)) bad |= MERCURY_ST_NONFINITE;
        acc += (double)x;
    }
    if (bad) atomicOr(a.status, bad);
    acc = warp_sum(acc);
    if (!first) wsync();
    if (lane == 0) sm.dred[0][warp] = acc;
    wsync();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int i = 0; i < kStepNT / 32; ++i) t += sm.dred[0][i];
        const unsigned long long bits = (unsigned long long)__double_as_longlong(t);
        *reinterpret_cast<uint4*>(a.sums + c) = make_uint4((uint32_t)bits, (uint32_t)(bits >> 32), (uint32_t)epoch, (uint32_t)(epoch >> 32));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// phase C: the walk.  One warp walks, and it must stay CONVERGED: every statement below is executed by all 32 lanes
// (stores of a warp-uniform value by every lane instead of `if (lane == 0)`), because a warp that has split pays
// hundreds of cycles for every shuffle and vote.  What the rare general paths need lives in shared memory (WalkEnv), so
// that the common path keeps its state in registers.
struct WalkEnv {
    WeightSrc src; int64_t n; float shift, totalw;
    const unsigned long long* pc; const uint32_t* ce; const uint32_t* qd0; const uint32_t* qd1; const uint32_t* stopw;
    const uint8_t* slot_of; uint32_t* sb; uint32_t* stage;
    const Pack2* slots; Pack2* spare; const Pack2* packs_g; const Summary* sums; uint32_t* start_bits;
    int head_k; uint32_t head_bits;   // chunk 0: the first head_k sub-chunks are already summed, the sum is head_bits
    unsigned int cnt_iter, cnt_serial, cnt_map;   // developer aid
};

// the reference's loop over 32 consecutive candidates held in shared memory, run redundantly by every lane
__device__ __forceinline__ uint32_t serial32(uint32_t b, uint32_t saddr) {
    uint4 q[kSub2 / 4];
#pragma unroll
    for (int i = 0; i < kSub2 / 4; ++i) q[i] = lds_u32x4(saddr + 16u * i);
    float s = b2f(b);
#pragma unroll
    for (int i = 0; i < kSub2 / 4; ++i) {
        s = __fadd_rn(s, b2f(q[i].x));
        s = __fadd_rn(s, b2f(q[i].y));
        s = __fadd_rn(s, b2f(q[i].z));
        s = __fadd_rn(s, b2f(q[i].w));
    }
    return f2b(s);
}

// candidates of sub-chunk f of chunk c that the pack does not carry: from the scores, into the stage line
__device__ __noinline__ uint32_t stage_sub2(const WalkEnv* env, int c, int f, int lane) {
    const int64_t i0 = (int64_t)c * kChunk + f * kSub2 + lane;
    float shift, total;
    shift = env->shift; total = env->totalw;
    env->stage[lane] = i0 < env->n ? f2b(env->src.get(i0, shift, total)) : 0u;
    __syncwarp();
    return smem_u32(env->stage);
}

// serial pass over sub-chunk f of chunk c (ef: its word in the pack): raw candidates from the pack (shared memory, address
// P) if it carries them
__device__ __forceinline__ uint32_t serial_sub2(const WalkEnv* env, uint32_t P, int c, int f, uint32_t ef, uint32_t b, int lane) {
    const uint32_t r1 = (ef >> 10) & 15u;   // raw slot + 1
    const uint32_t saddr = r1 ? P + (uint32_t)offsetof(Pack2, raw) + (r1 - 1u) * (kSub2 * 4u) : stage_sub2(env, c, f, lane);
    b = serial32(b, saddr);
    __syncwarp();
    return b;
}

// a flagged chunk, from sub-chunk k0 on (b: the exact sum in front of it).  P: the pack's address in SHARED memory.
// Expected case (the sum is in the chunk's guessed phase and changes phase once): the pack's prefix sums locate the
// sub-chunk whose end leaves the phase, that sub-chunk is summed with the reference's own loop (32 dependent fp32 adds),
// and the prefix sums for the next phase cross the rest.  Anything else: the general loop below -- the 64 increments
// for the current phase are scanned by the warp, the first sub-chunk that leaves the phase (or has no plain increment)
// is treated on its own (a 2-state map if it only has exact ties, else the serial loop), and the scan resumes behind it.
__device__ __forceinline__ uint32_t run_pack2(const WalkEnv* env, uint32_t P, int c, uint32_t b, int lane, int k0) {
    uint32_t E = phase_of(b), S = b - (E << 23);
    int k = k0;                                                      // first sub-chunk not yet applied
    const uint32_t e_lo = lds_u32(P + (uint32_t)offsetof(Pack2, e_lo));
    if (E == e_lo) {
        const uint2 pa = lds_u32x2(P + (uint32_t)offsetof(Pack2, psA) + 8u * lane);
        const uint32_t base = k0 > 0 ? lds_u32(P + (uint32_t)offsetof(Pack2, psA) + 4u * (k0 - 1)) : 0u;
        const uint32_t s_mid = S + (pa.x - base), s_out = S + (pa.y - base);   // (meaningful for sub-chunks >= k0)
        const unsigned over0 = __ballot_sync(0xffffffffu, 2 * lane >= k0 && s_mid >= kPhaseLimit);
        const unsigned over1 = __ballot_sync(0xffffffffu, 2 * lane + 1 >= k0 && s_out >= kPhaseLimit);
        if ((over0 | over1) == 0u) return bits_of(E, __shfl_sync(0xffffffffu, s_out, 31));
        const int f0 = over0 ? 2 * (__ffs(over0) - 1) : kSubs2, f1 = over1 ? 2 * (__ffs(over1) - 1) + 1 : kSubs2;
        const int f = min(f0, f1), fl = f >> 1;
        // the sum in front of sub-chunk f
        const uint32_t prev_out = __shfl_sync(0xffffffffu, s_out, fl > 0 ? fl - 1 : 0), this_mid = __shfl_sync(0xffffffffu, s_mid, fl);
        const uint32_t Sf = f == k0 ? S : ((f & 1) ? this_mid : prev_out);
        const uint32_t ef = lds_u32(P + (uint32_t)offsetof(Pack2, e) + 4u * f);
        const uint2 qa = lds_u32x2(P + (uint32_t)offsetof(Pack2, q0) + 8u * f), qb = lds_u32x2(P + (uint32_t)offsetof(Pack2, q1) + 8u * f);
        const uint32_t rest_lo = lds_u32(P + (uint32_t)offsetof(Pack2, psB) + 4u * f), rest_hi = lds_u32(P + (uint32_t)offsetof(Pack2, psB) + 4u * 63);
        const uint32_t g = ef & kPhaseMask;
        const uint32_t mx = g == E ? qa.x : qb.x, my = g == E ? qa.y : qb.y;
        const uint32_t Sn = Sf + ((Sf & 1u) ? my : mx);
        const bool by_map = g != kAnyPhase && (g == E || g + 1u == E) && mx != my && Sn < kPhaseLimit;
        k = f + 1;
        if (by_map) S = Sn;   // (exact ties inside, the sum stays in the phase: on with the general loop)
        else {
            b = serial_sub2(env, P, c, f, ef, bits_of(E, Sf), lane);
            E = phase_of(b); S = b - (E << 23);
            const uint32_t Se = S + (rest_hi - rest_lo);
            if (E == e_lo + 1u && Se < kPhaseLimit) return bits_of(E, Se);
        }
        if (k >= kSubs2) return bits_of(E, S);
    }
    const uint2 ee = lds_u32x2(P + (uint32_t)offsetof(Pack2, e) + 8u * lane);
    const uint4 q0 = lds_u32x4(P + (uint32_t)offsetof(Pack2, q0) + 16u * lane);  // maps of sub-chunks 2*lane, 2*lane+1 in their guessed phase
    const uint4 q1 = lds_u32x4(P + (uint32_t)offsetof(Pack2, q1) + 16u * lane);  // ... and in the phase above
    bool rescan = true;
    uint32_t s_in = 0u, s_mid = 0u, s_out = 0u;
    for (;;) {
        if (rescan) {
            const uint32_t w0 = 2 * lane >= k ? sub_inc(ee.x, q0.x, q0.y, q1.x, q1.y, E) : 0u;
            const uint32_t w1 = 2 * lane + 1 >= k ? sub_inc(ee.y, q0.z, q0.w, q1.z, q1.w, E) : 0u;
            uint32_t v = w0 + w1;                                    // (64 * 2^25 fits in 32 bits)
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, v, off);
                if (lane >= off) v += o;
            }
            s_in = S + (v - (w0 + w1));                              // running sum before sub-chunk 2*lane
            s_mid = s_in + w0; s_out = s_mid + w1;
        }
        // (sub-chunks before k were applied already: they never count as leaving the phase)
        const unsigned over0 = __ballot_sync(0xffffffffu, 2 * lane >= k && s_mid >= kPhaseLimit);
        const unsigned over1 = __ballot_sync(0xffffffffu, 2 * lane + 1 >= k && s_out >= kPhaseLimit);
        if ((over0 | over1) == 0u) { S = __shfl_sync(0xffffffffu, s_out, 31); break; }
        const int f0 = over0 ? 2 * (__ffs(over0) - 1) : kSubs2, f1 = over1 ? 2 * (__ffs(over1) - 1) + 1 : kSubs2;
        const int f = min(f0, f1);
        const int fl = f >> 1;
        const bool odd = f & 1;
        const uint32_t Sf = __shfl_sync(0xffffffffu, odd ? s_mid : s_in, fl);
        const uint32_t ef = __shfl_sync(0xffffffffu, odd ? ee.y : ee.x, fl);
        const uint32_t qa = __shfl_sync(0xffffffffu, odd ? q0.z : q0.x, fl), qb = __shfl_sync(0xffffffffu, odd ? q0.w : q0.y, fl);
        const uint32_t qc = __shfl_sync(0xffffffffu, odd ? q1.z : q1.x, fl), qd = __shfl_sync(0xffffffffu, odd ? q1.w : q1.y, fl);
        const uint32_t g = ef & kPhaseMask;
        // exact ties inside: the 2-state map, if the sum stays in the phase
        const uint32_t mx = g == E ? qa : qc, my = g == E ? qb : qd;
        const uint32_t Sn = Sf + ((Sf & 1u) ? my : mx);
        const bool by_map = g != kAnyPhase && (g == E || g + 1u == E) && mx != my && Sn < kPhaseLimit;
        if (by_map) {
            // same phase: every running sum behind f moves by (what f really added) - (the kBadInc it was counted as);
            // no new scan (unsigned wrap-around is fine: only entries behind f are looked at again)
            const uint32_t adj = (Sn - Sf) - kBadInc;
            S = Sn;
            s_in += adj; s_mid += adj; s_out += adj;
            rescan = false;
        } else {
            b = serial_sub2(env, P, c, f, ef, bits_of(E, Sf), lane);
            E = phase_of(b); S = b - (E << 23);
            rescan = true;
        }
        k = f + 1;
        if (k >= kSubs2) break;
    }
    return bits_of(E, S);
}

// a pack that is not among the prefetched ones: copied into the spare slot; returns its shared address
__device__ __noinline__ uint32_t fetch_pack2(const WalkEnv* env, int slot_global, int lane) {
    const uint4* g4 = reinterpret_cast<const uint4*>(env->packs_g + slot_global);
    uint4* s4 = reinterpret_cast<uint4*>(env->spare);
    for (int i = lane; i < (int)(sizeof(Pack2) / 16); i += 32) s4[i] = __ldcg(g4 + i);
    __syncwarp();
    return smem_u32(env->spare);
}

// one whole chunk summed the general way (serial_chunk of the kernel chain); only reached when a guess was contradicted
__device__ __noinline__ uint32_t slow_chunk2(const WalkEnv* env, int c, uint32_t b, int lane) {
    SpineCtx x;
    x.src = env->src; x.n = env->n; x.shift = env->shift; x.totalw = env->totalw; x.lane = lane; x.dbg = nullptr; x.dbgt = nullptr;
    x.prof_ = false; x.pacc = nullptr; x.stage = env->stage;
    return serial_chunk<false>(x, c, b);
}

// chunks [c, c_end) the general way (no precompiled step, or its guard was contradicted); records their boundary values
__device__ __noinline__ uint32_t walk_general2(const WalkEnv* env, int c, int c_end, uint32_t b, int lane) {
    while (c < c_end) {
        const uint32_t E = phase_of(b), S = b - (E << 23);
        const uint32_t ec = env->ce[c], g = ec & kPhaseMask;
        env->start_bits[c] = b; env->sb[c] = b;   // (every lane stores the same word)
        if (ec & kFlag) {
            const int sl = (int)env->slot_of[c] - 1;
            const uint32_t P = sl >= 0 ? smem_u32(&env->slots[sl]) : fetch_pack2(env, (int)env->sums[c].pad[0], lane);
            const bool head = c == 0 && env->head_k > 0;
            b = run_pack2(env, P, c, head ? env->head_bits : b, lane, head ? env->head_k : 0);
            __syncwarp();
            ++c;
            continue;
        }
        if (g == kAnyPhase) { ++c; continue; }
        if (g != E || (ec & kNoPack)) { b = slow_chunk2(env, c, b, lane); ++c; continue; }
        if (ec & kTie0) {
            const uint32_t Sn = apply(Pair{env->qd0[c], env->qd1[c]}, S);
            if (Sn >= kPhaseLimit) b = slow_chunk2(env, c, b, lane); else b = bits_of(E, Sn);
            ++c;
            continue;
        }
        const int s = min(next_bit(env->stopw, c + 1), c_end);
        const unsigned long long base = env->pc[c];
        const unsigned long long Sn = (unsigned long long)S + (env->pc[s] - base);
        int end = s;
        if (Sn >= kPhaseLimit) {
            int lo = c, hi = s;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if ((unsigned long long)S + (env->pc[mid] - base) >= kPhaseLimit) hi = mid; else lo = mid;
            }
            end = lo;
        }
        for (int k = c + 1 + lane; k < end; k += 32) {
            const uint32_t v = bits_of(E, S + (uint32_t)(env->pc[k] - base));
            env->start_bits[k] = v; env->sb[k] = v;
        }
        __syncwarp();
        if (end == s) { b = bits_of(E, (uint32_t)Sn); c = s; continue; }
        const uint32_t bf = bits_of(E, S + (uint32_t)(env->pc[end] - base));
        env->start_bits[end] = bf; env->sb[end] = bf;
        b = slow_chunk2(env, end, bf, lane);
        c = end + 1;
    }
    return b;
}

__host__ __device__ inline size_t walk2_fixed_bytes(int nchunks) {
    size_t b = (size_t)(nchunks + 1) * 8 + (size_t)nchunks * 13 + (size_t)((nchunks + 32) / 32) * 4 + (size_t)(nchunks + 1) * 4;
    b = (b + 127) & ~(size_t)127;
    return b + (size_t)kStageWords * 4 + (size_t)kStepMaxEvents * 20;
}
__host__ __device__ inline int walk2_slots(int nchunks) {
    const long fit = ((long)kStepSmemBytes - (long)walk2_fixed_bytes(nchunks) - 256) / (long)sizeof(Pack2) - 1;
    return (int)(fit > 16 ? 16 : fit);
}

struct WalkStatic {
    WalkEnv env;
    int npacks;
    bool packs_early;
    unsigned long long wtot[kStepMaxChunks / kStepNT][kStepNT / 32];
    int wcnt[kStepMaxChunks / kStepNT][kStepNT / 32];
};

// fp64 mode, phase C (the walker CTA): scan of the chunk sums, the chunk of every draw, publish.  No serial part.
__device__ __forceinline__ void walker_f64_body(const StepArgs& a, unsigned char* sm, WalkStatic& ws, float stat_ema, unsigned long long epoch) {
    const int nchunks = a.nchunks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* ap = reinterpret_cast<double*>(sm);   // [nchunks + 1]
    constexpr int kR = kStepMaxChunks / kStepNT;
    const int nr = (nchunks + kStepNT - 1) / kStepNT;
    double v[kR], excl[kR];
    {
        const uint32_t tag_lo = (uint32_t)epoch, tag_hi = (uint32_t)(epoch >> 32);
        const unsigned long long t0 = global_ns();
        bool ok[kR];
#pragma unroll
        for (int r = 0; r < kR; ++r) { ok[r] = !(r < nr && r * kStepNT + tid < nchunks); v[r] = 0.0; }
        for (unsigned int it = 0;; ++it) {
            bool all = true;
#pragma unroll
            for (int r = 0; r < kR; ++r) {
                if (!ok[r]) {
                    uint4 q;
                    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
                                 : "l"(a.sums + (r * kStepNT + tid)) : "memory");
                    if (q.z == tag_lo && q.w == tag_hi) { v[r] = __longlong_as_double((long long)(((unsigned long long)q.y << 32) | q.x)); ok[r] = true; }
                    else all = false;
                }
            }
            if (all) break;
            if ((it & 15u) == 15u && global_ns() - t0 > kStepSpinNs) { atomicOr(a.status, MERCURY_ST_INTERNAL); break; }
        }
    }
    double* dw = reinterpret_cast<double*>(&ws.wtot[0][0]);   // [kR][8] warp totals
#pragma unroll
    for (int r = 0; r < kR; ++r) {
        if (r >= nr) break;
        double inc = v[r];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) dw[r * (kStepNT / 32) + warp] = inc;
        excl[r] = inc - v[r];
    }
    wsync();
    {
        double running = 0.0;
#pragma unroll
        for (int r = 0; r < kR; ++r) {
            if (r >= nr) break;
            const int c = r * kStepNT + tid;
            double pre = running;
#pragma unroll
            for (int i = 0; i < kStepNT / 32; ++i) {
                const double t = dw[r * (kStepNT / 32) + i];
                if (i < warp) pre += t;
                running += t;
            }
            if (c < nchunks) { ap[c] = pre + excl[r]; a.approx[c] = pre + excl[r]; }
        }
        if (tid == 0) {
            ap[nchunks] = running; a.approx[nchunks] = running;
            *a.ema_state = stat_ema;
            a.stats[MERCURY_STAT_CHAIN_TOTAL] = (float)running;
            if (!(running > 0.0)) atomicOr(a.status, MERCURY_ST_ZERO_SUM);
            else if (isinf((float)running)) atomicOr(a.status, MERCURY_ST_NONFINITE);
            a.ctrl->times[6] = global_ns();
        }
    }
    wsync();
    const double total64 = ap[nchunks];
    for (int64_t j = tid; j < a.d.b; j += kStepNT) {
        const double uj = a.d.u[j];
        int lo = 0, hi = nchunks - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((double)(float)(ap[mid + 1] / total64) < uj) lo = mid + 1; else hi = mid;
        }
        a.ctrl->dapprox[j] = ap[lo];
        a.ctrl->dtotal[j] = total64;
        __threadfence();
        st_relaxed_u64(&a.ctrl->dinfo[j], (1ull << 63) | ((unsigned long long)lo << 32));
        st_relaxed_u64(&a.ctrl->dinfo2[j], 1ull << 63);
    }
    if (tid == 0) a.ctrl->times[7] = global_ns();
}

// The walker (one CTA, 256 threads + named barrier 1).  Leaves the boundary values in global memory, and publishes the
// chunk of every draw.
__device__ __forceinline__ void walker_body(const StepArgs& a, unsigned char* sm, WalkStatic& ws, float stat_shift, float stat_total,
                                            float stat_ema, unsigned long long epoch) {
    const int nchunks = a.nchunks;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long* pc = reinterpret_cast<unsigned long long*>(sm);
    uint32_t* ce = reinterpret_cast<uint32_t*>(pc + nchunks + 1);
    uint32_t* qd0 = ce + nchunks; uint32_t* qd1 = qd0 + nchunks;
    uint32_t* stopw = qd1 + nchunks;
    uint32_t* sb = stopw + (nchunks + 32) / 32;                        // [nchunks+1] boundary values
    uint8_t* slot_of = reinterpret_cast<uint8_t*>(sb + nchunks + 1);
    size_t off = (size_t)(nchunks + 1) * 8 + (size_t)nchunks * 13 + (size_t)((nchunks + 32) / 32) * 4 + (size_t)(nchunks + 1) * 4;
    off = (off + 127) & ~(size_t)127;
    const int max_slots = walk2_slots(nchunks);
    Pack2* slots = reinterpret_cast<Pack2*>(sm + off);
    Pack2* spare = slots + max_slots;
    uint32_t* stage = reinterpret_cast<uint32_t*>(sm + off + (size_t)(max_slots + 1) * sizeof(Pack2));
    uint4* evs = reinterpret_cast<uint4*>(stage + kStageWords);
    uint32_t* evb = reinterpret_cast<uint32_t*>(evs + kStepMaxEvents);

    // ---- the summaries (the packs follow once the steps are compiled: they are still being written)
    if (tid == 0) {
        WalkEnv& e = ws.env;   // (head_k / head_bits were filled in by the warp that summed the head of the pool)
        e.src = a.d.src; e.n = a.n; e.shift = stat_shift; e.totalw = stat_total;
        e.pc = pc; e.ce = ce; e.qd0 = qd0; e.qd1 = qd1; e.stopw = stopw; e.slot_of = slot_of; e.sb = sb; e.stage = stage;
        e.slots = slots; e.spare = spare; e.packs_g = a.packs; e.sums = a.sums; e.start_bits = a.start_bits;
        e.cnt_iter = e.cnt_serial = e.cnt_map = 0u;
        a.ctrl->times[13] = (unsigned long long)e.head_k;
    }
    // All rounds of 256 chunks are handled together, so that every stage below needs ONE barrier: loads, prefix sums of the
    // chunk increments, the stop mask, the precompiled steps.
    constexpr int kR = kStepMaxChunks / kStepNT;
    const int nr = (nchunks + kStepNT - 1) / kStepNT;
    uint4 lo[kR];
    uint32_t sl[kR];
    {
        // every thread polls its own summaries until they carry this launch's tag (the workers write them with two 16-byte
        // stores, the tag in both: whatever is read with the tag is complete)
        const uint32_t tag_lo = (uint32_t)epoch, tag_hi = (uint32_t)(epoch >> 32);
        const unsigned long long t0 = global_ns();
        bool ok[kR];
#pragma unroll
        for (int r = 0; r < kR; ++r) { ok[r] = !(r < nr && r * kStepNT + tid < nchunks); lo[r] = make_uint4(0u, 0u, 0u, 0u); sl[r] = 0u; }
        for (unsigned int it = 0;; ++it) {
            bool all = true;
#pragma unroll
            for (int r = 0; r < kR; ++r) {
                if (!ok[r]) {
                    const uint32_t* src = reinterpret_cast<const uint32_t*>(a.sums + (r * kStepNT + tid));
                    uint4 v; uint2 h;
                    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src) : "memory");
                    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0,%1}, [%2];" : "=r"(h.x), "=r"(h.y) : "l"(src + 4) : "memory");
                    if (v.w == tag_lo && h.x == tag_hi) { lo[r] = v; sl[r] = h.y; ok[r] = true; } else all = false;
                }
            }
            if (all) break;
            if ((it & 15u) == 15u && global_ns() - t0 > kStepSpinNs) { atomicOr(a.status, MERCURY_ST_INTERNAL); break; }
        }
    }
    // the packs: all slots are handed out by now (a flagged chunk claims its slot before it writes its summary, and every
    // summary has been seen); if the packs are all written already they are fetched now, else once the steps are compiled
    wsync();
    if (tid == 64) a.ctrl->times[4] = global_ns();
    if (tid == 0) {
        const unsigned int nflag = __ldcg(&a.ctrl->nflag);
        const unsigned int want = nflag < (unsigned)kStepMaxPacks ? nflag : (unsigned)kStepMaxPacks;
        unsigned int got;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(got) : "l"(&a.ctrl->packs_done) : "memory");
        ws.npacks = (int)want;
        ws.packs_early = got >= want;
        if (got >= want) __threadfence();
    }
    wsync();
    auto load_packs = [&]() {
        constexpr int kPackVecs = (int)(sizeof(Pack2) / 16);
        const uint4* g4 = reinterpret_cast<const uint4*>(a.packs);
        uint4* s4 = reinterpret_cast<uint4*>(slots);
        const int total_vecs = min(max_slots, ws.npacks) * kPackVecs;
        for (int i0 = tid; i0 < total_vecs; i0 += 8 * kStepNT) {
            uint4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) { const int i = i0 + k * kStepNT; if (i < total_vecs) v[k] = __ldcg(g4 + i); }
#pragma unroll
            for (int k = 0; k < 8; ++k) { const int i = i0 + k * kStepNT; if (i < total_vecs) s4[i] = v[k]; }
        }
    };
    const bool packs_early = ws.packs_early;
    if (packs_early) load_packs();
    unsigned long long excl[kR];   // this chunk's increment prefix inside its warp (exclusive)
#pragma unroll
    for (int r = 0; r < kR; ++r) {
        if (r >= nr) break;
        const int c = r * kStepNT + tid;
        unsigned long long inc = 0ull;
        if (c < nchunks) {
            ce[c] = lo[r].x; qd0[c] = lo[r].y; qd1[c] = lo[r].z;
            slot_of[c] = ((lo[r].x & kFlag) && (int)sl[r] < max_slots) ? (uint8_t)(sl[r] + 1) : 0;
            inc = lo[r].y;
        }
        const unsigned long long mine = inc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) ws.wtot[r][warp] = inc;
        excl[r] = inc - mine;
    }
    wsync();
    {
        unsigned long long running = 0ull;
#pragma unroll
        for (int r = 0; r < kR; ++r) {
            if (r >= nr) break;
            const int c = r * kStepNT + tid;
            unsigned long long pre = running;
#pragma unroll
            for (int i = 0; i < kStepNT / 32; ++i) {
                const unsigned long long t = ws.wtot[r][i];
                if (i < warp) pre += t;
                running += t;
            }
            if (c < nchunks) pc[c] = pre + excl[r];
        }
        if (tid == 0) pc[nchunks] = running;
    }
    // stop mask: chunks that need individual treatment, or where the guess changes
    for (int w0 = 0; w0 < (nchunks + 32) / 32; w0 += kStepNT / 32) {
        const int c = (w0 + warp) * 32 + lane;
        bool stop = false;
        if (c < nchunks) {
            const uint32_t e = ce[c], g = e & kPhaseMask;
            stop = (e & (kFlag | kTie0 | kNoPack)) != 0u;
            if (c > 0) {   // (a change of the guess right behind a chunk of its own needs no stop: a run starts there anyway)
                const uint32_t ep = ce[c - 1], gp = ep & kPhaseMask;
                if (g != kAnyPhase && gp != kAnyPhase && g != gp && !(ep & (kFlag | kTie0 | kNoPack))) stop = true;
            }
        }
        if (c == nchunks) stop = true;  // sentinel
        const unsigned bits = __ballot_sync(0xffffffffu, stop);
        if (lane == 0 && w0 + warp < (nchunks + 32) / 32) stopw[w0 + warp] = bits;
    }
    wsync();
    // The walk, precompiled: one step per stop s (and one for the end of the pool).  A step is the additive run of plain
    // chunks in front of s -- one guarded addition of prefix sums -- followed by what chunk s itself needs: its pack, its
    // 2-state map, the general code, or nothing (s only marks a change of the guessed phase).
    unsigned bal[kR];
#pragma unroll
    for (int r = 0; r < kR; ++r) {
        bal[r] = 0u;
        if (r >= nr) break;
        const int c = r * kStepNT + tid;
        const bool stop = c < nchunks && ((stopw[c >> 5] >> (c & 31)) & 1u);
        bal[r] = __ballot_sync(0xffffffffu, stop);
        if (lane == 0) ws.wcnt[r][warp] = __popc(bal[r]);
    }
    wsync();
    int nev_all = 0;
    auto make_step = [&](int s2) -> uint4 {
        // the run in front of s2 starts behind the previous stop if that was a chunk of its own, else at it
        int p = s2 - 1, run_start = 0;
        {
            int w = p >> 5;
            uint32_t m = p >= 0 ? stopw[w] & (0xffffffffu >> (31 - (p & 31))) : 0u;
            while (p >= 0 && m == 0u && w > 0) m = stopw[--w];
            if (p >= 0 && m != 0u) {
                const int q = (w << 5) + 31 - __clz(m);
                run_start = (ce[q] & (kFlag | kTie0 | kNoPack)) ? q + 1 : q;
            }
        }
        uint32_t g = kAnyPhase;
        for (int k = run_start; k < s2 && g == kAnyPhase; ++k) g = ce[k] & kPhaseMask;
        const unsigned long long inc = pc[s2] - pc[run_start];
        uint32_t kind = kEvEnd;
        if (s2 < nchunks) {
            const uint32_t e = ce[s2];
            kind = (e & kFlag) ? kEvPack : (e & kNoPack) ? kEvSlow : (e & kTie0) ? kEvMap : kEvNone;
        }
        return make_uint4(kind | (g << 8), (uint32_t)run_start | ((uint32_t)s2 << 16), (uint32_t)inc, (uint32_t)(inc >> 32));
    };
#pragma unroll
    for (int r = 0; r < kR; ++r) {
        if (r >= nr) break;
        const int c = r * kStepNT + tid;
        int idx = nev_all + __popc(bal[r] & ((1u << lane) - 1u));
#pragma unroll
        for (int i = 0; i < kStepNT / 32; ++i) {
            const int t = ws.wcnt[r][i];
            if (i < warp) idx += t;
            nev_all += t;
        }
        if (((bal[r] >> lane) & 1u) && idx < kStepMaxEvents) evs[idx] = make_step(c);
    }
    if (tid == 0 && nev_all < kStepMaxEvents) evs[nev_all] = make_step(nchunks);   // the end of the pool
    nev_all += 1;
    if (!packs_early) {
        if (tid == 0) {   // wait until every pack is written
            const unsigned int want = (unsigned int)ws.npacks;
            unsigned int got;
            const unsigned long long t0 = global_ns();
            for (unsigned int it = 0;; ++it) {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(got) : "l"(&a.ctrl->packs_done) : "memory");
                if (got >= want) break;
                if ((it & 63u) == 63u && global_ns() - t0 > kStepSpinNs) { atomicOr(a.status, MERCURY_ST_INTERNAL); break; }
            }
            __threadfence();
        }
        wsync();
        load_packs();
    }
    if (tid == 64) a.ctrl->times[10] = global_ns();
    wsync();
    const int nev = nev_all;
    const bool use_events = nev <= kStepMaxEvents;
    if (tid == 32) a.ctrl->times[5] = global_ns();   // (a thread of another warp: the walking warp stays converged)

    const int walk_warp = (a.dbg_flags >> 4) & 7;    // developer aid: which warp walks (default 0)
    if (warp == walk_warp) {
        const WalkEnv* env = &ws.env;
        const bool stamp = a.dbg_flags & 1;
        __syncwarp();
        if (a.dbg_flags & 2) {   // developer aid: how fast does this warp run right now?
            long long c0 = clock64();
            uint32_t xx = (uint32_t)c0;
#pragma unroll 1
            for (int r = 0; r < 16; ++r) xx += __shfl_up_sync(0xffffffffu, xx, 1);
            long long c1 = clock64();
            float ff = __uint_as_float(xx & 0x3fffffffu);
#pragma unroll
            for (int r = 0; r < 64; ++r) ff = __fadd_rn(ff, 1.5f);
            long long c2 = clock64();
            a.ctrl->evt[104] = c1 - c0; a.ctrl->evt[105] = c2 - c1; a.ctrl->evt[108] = xx + __float_as_uint(ff);
        }
        uint32_t b = 0u;
        if (!use_events) {
            b = walk_general2(env, 0, nchunks, b, lane);
        } else {
            uint32_t E = 0u, S = 0u;
            if (stamp) a.ctrl->evt[0] = (unsigned long long)clock64() << 16;
            for (int i = 0; i < nev; ++i) {
                const uint4 ev = evs[i];
                if (stamp && i > 0 && i < 100) a.ctrl->evt[i] = ((unsigned long long)clock64() << 16) | (evs[i - 1].x & 0xfu);
                const uint32_t kind = ev.x & 0xffu, ph = (ev.x >> 8) & kPhaseMask;
                const int c = (int)(ev.y & 0xffffu), s = (int)(ev.y >> 16);   // the run [c, s), then chunk s
                if (s > c) {
                    const unsigned long long Sn = (unsigned long long)S + (((unsigned long long)ev.w << 32) | ev.z);
                    if ((ph == E || ph == kAnyPhase) && Sn < kPhaseLimit) {
                        evb[i] = bits_of(E, S);   // boundary values of [c, s): filled in by the whole CTA afterwards
                        S = (uint32_t)Sn;
                    } else {   // a contradicted guard: the general walk over the run (it records the boundary values itself)
                        evs[i].x = ev.x | 0x80u;
                        __syncwarp();
                        b = walk_general2(env, c, s, bits_of(E, S), lane);
                        E = phase_of(b); S = b - (E << 23);
                    }
                }
                if (kind == kEvPack) {
                    b = bits_of(E, S);
                    a.start_bits[s] = b; sb[s] = b;
                    const int sl = (int)slot_of[s] - 1;
                    const uint32_t P = sl >= 0 ? smem_u32(&slots[sl]) : fetch_pack2(env, (int)a.sums[s].pad[0], lane);
                    const bool head = s == 0 && env->head_k > 0;
                    b = run_pack2(env, P, s, head ? env->head_bits : b, lane, head ? env->head_k : 0);
                    __syncwarp();
                    E = phase_of(b); S = b - (E << 23);
                } else if (kind == kEvMap) {
                    const uint32_t d0 = qd0[s], d1 = qd1[s];
                    const uint32_t Sn = S + ((S & 1u) ? d1 : d0);
                    if ((ce[s] & kPhaseMask) == E && Sn < kPhaseLimit) {
                        a.start_bits[s] = bits_of(E, S); sb[s] = bits_of(E, S);
                        S = Sn;
                    } else {
                        b = walk_general2(env, s, s + 1, bits_of(E, S), lane);
                        E = phase_of(b); S = b - (E << 23);
                    }
                } else if (kind == kEvSlow) {
                    b = walk_general2(env, s, s + 1, bits_of(E, S), lane);
                    E = phase_of(b); S = b - (E << 23);
                }
            }
            b = bits_of(E, S);
            if (stamp && nev < 100) a.ctrl->evt[nev] = ((unsigned long long)clock64() << 16) | (evs[nev - 1].x & 0xfu);
        }
        // (every lane writes the same words)
        *a.ema_state = stat_ema;  // EMAverage.value for the next call (util.py:207-213)
        a.start_bits[nchunks] = b; sb[nchunks] = b;
        const float total = b2f(b);
        *a.total = total;
        a.stats[MERCURY_STAT_CHAIN_TOTAL] = total;
        if (lane == 0) {
            if (!(total > 0.f)) atomicOr(a.status, MERCURY_ST_ZERO_SUM);
            else if (isinf(total)) atomicOr(a.status, MERCURY_ST_NONFINITE);
            a.ctrl->times[6] = global_ns();
            a.ctrl->times[12] = (unsigned long long)nev;
        }
    }
    wsync();
    // ---- boundary values inside the additive steps (whole CTA, one warp per step)
    if (use_events) {
        for (int i = warp; i < nev; i += kStepNT / 32) {
            const uint4 ev = evs[i];
            if (ev.x & 0x80u) continue;   // (that run was walked the general way)
            const int c = (int)(ev.y & 0xffffu), s = (int)(ev.y >> 16);
            if (s <= c) continue;
            const uint32_t b0 = evb[i], E = phase_of(b0), S = b0 - (E << 23);
            const unsigned long long base = pc[c];
            for (int k = c + lane; k < s; k += 32) {
                const uint32_t v = bits_of(E, S + (uint32_t)(pc[k] - base));
                a.start_bits[k] = v; sb[k] = v;
            }
        }
    }
    wsync();
    // ---- sharded pool: every rank's chain total, exchanged through the peers' boxes (one CTA waits here, not every draw)
    if (a.peers) {
        float* sh_cdf = reinterpret_cast<float*>(evs);   // [nshards] (the steps are not needed any more)
        PeerBox* const mine = a.peers[a.rank];
        const unsigned int tag = mine->seq_total + 1u;
        if (tid < a.nshards) {
            peer_store(&a.peers[tid]->total[tag & 1u][a.rank], ((unsigned long long)sb[nchunks] << 32) | tag);
            unsigned long long w;
            if (!peer_wait(&mine->total[tag & 1u][tid], tag, w, mine)) atomicOr(a.status, MERCURY_ST_PEER_TIMEOUT);
            sh_cdf[tid] = __uint_as_float((unsigned int)(w >> 32));
            a.totals_all[tid] = sh_cdf[tid];
        }
        wsync();
        if (tid == 0) {   // serial fp32 CDF over the shard totals (shard_select_kernel's arithmetic); step counters for the next call
            mine->seq_total = tag; mine->seq_sum = mine->seq_sum + 1u;
            float cs = 0.f;
            for (int k = 0; k < a.nshards; ++k) cs = __fadd_rn(cs, sh_cdf[k]);
            float run = 0.f;
            for (int k = 0; k < a.nshards; ++k) { run = __fadd_rn(run, sh_cdf[k]); sh_cdf[k] = __fdiv_rn(run, cs); }
        }
        wsync();
    }
    // ---- the chunk of every draw: first chunk whose last cdf value is not below u (ATen's lower bound), then publish
    const float total = b2f(sb[nchunks]);
    const bool poisoned = a.peers && (*reinterpret_cast<volatile int32_t*>(a.status) & MERCURY_ST_PEER_TIMEOUT) != 0;
    for (int64_t j = tid; j < a.d.b; j += kStepNT) {
        double uj = a.d.u[j];
        if (a.peers) {   // which shard the uniform falls in, and where inside that shard's own CDF (-1: another rank's draw)
            const float* sh_cdf = reinterpret_cast<const float*>(evs);
            int k = 0;
            while (k < a.nshards - 1 && (double)sh_cdf[k] < uj) ++k;
            double lu = -1.0;
            if (k == a.rank) {
                const double lo_c = k > 0 ? (double)sh_cdf[k - 1] : 0.0, hi_c = (double)sh_cdf[k];
                lu = hi_c > lo_c ? (uj - lo_c) / (hi_c - lo_c) : 0.0;
                if (lu < 0.0) lu = 0.0;
                if (lu >= 1.0) lu = 0.99999999999999989;
            }
            // (the draw's CTA takes the rescaled uniform from dinfo3: a word that is valid on its own, so no fence is needed)
            if (poisoned) {   // a peer exchange of this step timed out (see peer_poison)
                a.local_u[j] = peer_poison();
                st_relaxed_u64(&a.ctrl->dinfo3[j], (1ull << 63) | (unsigned long long)__double_as_longlong(peer_poison()));
                st_relaxed_u64(&a.ctrl->dinfo[j], 1ull << 63);
                st_relaxed_u64(&a.ctrl->dinfo2[j], (1ull << 63) | f2b(total));
                continue;
            }
            a.local_u[j] = lu;
            if (lu < 0.0) {
                st_relaxed_u64(&a.ctrl->dinfo[j], (1ull << 63) | (1ull << 62));
                st_relaxed_u64(&a.ctrl->dinfo2[j], (1ull << 63) | f2b(total));
                continue;
            }
            uj = lu;
            st_relaxed_u64(&a.ctrl->dinfo3[j], (1ull << 63) | (unsigned long long)__double_as_longlong(lu));   // (lu >= 0: bit 63 is free)
        }
        int lo = 0, hi = nchunks - 1;          // answer in [lo, hi]
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((double)__fdiv_rn(b2f(sb[mid + 1]), total) < uj) lo = mid + 1; else hi = mid;
        }
        st_relaxed_u64(&a.ctrl->dinfo[j], (1ull << 63) | ((unsigned long long)lo << 32) | sb[lo]);
        st_relaxed_u64(&a.ctrl->dinfo2[j], (1ull << 63) | f2b(total));
    }
    if (tid == 0) a.ctrl->times[7] = global_ns();
}

// Tiles of scores per CTA: contiguous ranges, weighted by thirds of the grid (with three CTAs per SM the grid comes up in
// three waves).  A static rule, so the sums come out the same in every run.  first tile of CTA b = T * W(b) / W(G), W =
// cumulative weight.
// (the three weights and the two split points -- 0: thirds of the grid -- live in constant memory so that the developer knob
// MERCURY_STEP_W="w1,w2,w3,g1,g2" can replace them without a rebuild)
// Equal weights since the launch carries the programmatic-serialization attribute: in back-to-back steps the CTAs are
// dispatched while the previous step drains and the waves no longer start a microsecond apart (53 : 50 : 47, tuned before
// that, measures 46.9 us per step against 46.3 us for 50 : 50 : 50 on the same box; more skew is worse either way).
__constant__ unsigned int c_tile_w[5] = {50u, 50u, 50u, 0u, 0u};
__device__ __forceinline__ unsigned int tile_weight_prefix(unsigned int b, unsigned int G) {
    const unsigned int g1 = c_tile_w[3] ? min(c_tile_w[3], G) : G / 3u, g2 = c_tile_w[4] ? min(max(c_tile_w[4], g1), G) : 2u * G / 3u;
    return c_tile_w[0] * min(b, g1) + c_tile_w[1] * (min(b, g2) - min(b, g1)) + c_tile_w[2] * (b - min(b, g2));
}
__device__ __forceinline__ unsigned int tile_begin(unsigned int b, unsigned int G, unsigned int T) {
    return (unsigned int)((unsigned long long)T * tile_weight_prefix(b, G) / tile_weight_prefix(G, G));
}

// The chunk a CTA's first draw will most likely land in, from the fp64 sums of the scores (the exact chain differs from them
// by its rounding drift, ~1e-5 relative: the estimate names the right chunk unless u sits that close to a chunk boundary --
// 99.4 % of the draws at cfg2), and that chunk's probabilities, fetched into cdf_s in draw_resolve's layout.  Runs while the
// walk is under way and the CTA has nothing else to do; if the walker's word names the same chunk, the resolve starts from
// shared memory, if not, nothing is lost.  Not inlined: its registers must not lengthen the live ranges of the phases
// around it (the kernel sits at its register limit).
__device__ __noinline__ int spec_chunk2(const double* __restrict__ approx, int nchunks, int64_t n, const float* __restrict__ scores,
                                        double u0, float shift, float total, double* dred1, float* cdf_s) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // approx[c]: the fp64 sum of the scores in front of chunk c, left there by the worker that owns the chunk (phase B0).  No
    // synchronisation with those writes is needed: a stale or missing value can only make the guess wrong, never the result.
    // (An estimate: the boundary's running total is compared with u * total -- no division per boundary.)
    const double thr = u0 * (double)total, sh_d = (double)shift * (double)kChunk;
    int cnt = 0;
    for (int c = tid + 1; c < nchunks; c += kStepNT)
        if (__ldcg(approx + c) + (double)c * sh_d < thr) ++cnt;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    wsync();   // (the reduction scratch is free: this CTA's summaries are out)
    if (lane == 0) dred1[warp] = (double)cnt;
    wsync();
    int c_est = 0;
#pragma unroll
    for (int i = 0; i < kStepNT / 32; ++i) c_est += (int)dred1[i];
    const int chat = min(c_est, nchunks - 1);
    const int64_t start = (int64_t)chat * kChunk;
    const int len = (int)min((int64_t)kChunk, n - start);
    const int base = tid * kDrawIPT;
#pragma unroll
    for (int k = 0; k < kDrawIPT; ++k)   // p = (score + shift) / total, WeightSrc::get's arithmetic
        cdf_s[base + k] = base + k < len ? __fdiv_rn(__fadd_rn(scores[start + base + k], shift), total) : 0.f;
    return chat;
}

// ---------------------------------------------------------------------------------------------------------------
// F64 = true: the non-exact mode as a kernel of its own, without the exact mode's code (29.5 -> 27.7 us per step).  The
// general kernel (F64 = false) still decides at run time and keeps both: taking the fp64 code OUT of it made the exact mode a
// microsecond slower (45.8 against 44.8 us, same box) -- at 250 KB of code and the register limit, the compiler's choices
// for the hot phases move with everything around them; measured, kept as it was.
template <typename T, int CT, bool F64>
__global__ void __launch_bounds__(kStepThreads, 3) fused_step_kernel(StepArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ Sum2Smem s2;
    __shared__ DrawSmem ds;
    __shared__ WalkStatic wst;
    __shared__ unsigned long long dinfo_s, dinfo3_s;
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_wait();                 // (no-ops unless launched with programmatic serialization: nothing of the previous launch on the
    pdl_launch_dependents();    //  stream is touched before it has completed; the next launch may be dispatched into free slots)
    if (blockIdx.x == 0 && tid == 0) {
        const unsigned long long now = global_ns();
        a.ctrl->nflag = 0u; a.ctrl->packs_done = 0u;
        a.ctrl->epoch = now;                       // this launch's tag (read by everybody behind the grid barrier)
        a.ctrl->times[11] = 0ull;
        a.ctrl->times[0] = now;
    }
    if (blockIdx.x == 0) for (int64_t j = tid; j < a.d.b; j += kStepThreads) { a.ctrl->dinfo[j] = 0ull; a.ctrl->dinfo2[j] = 0ull; }
    if (blockIdx.x == 0 && a.peers) {
        for (int64_t j = tid; j < a.d.b; j += kStepThreads) a.ctrl->dinfo3[j] = 0ull;
        if (tid < 2) a.ctrl->gsum[tid] = 0ull;
    }
    __shared__ unsigned int b0_s[2];
    // ---- A: scores + tile sums
    // (contiguous tiles per CTA, so that one partial sum per CTA gives the pool statistics and the running totals in front
    // of the chunks: 2 loads per thread behind the barrier instead of every CTA re-reading every tile's sum)
    {
        // (the CTA's tile range: two 64-bit divisions, by two threads instead of by everybody in front of the first load)
        __shared__ unsigned int range_s[2];
        if (tid < 2) range_s[tid] = tile_begin(blockIdx.x + (unsigned int)tid, gridDim.x, (unsigned int)a.ntiles);
        score_small_body<T, CT, kStepStages, kStepNT, true, false>(smem, static_cast<const T*>(a.logits), a.labels, a.n, a.c, a.tile_rows,
                                                                   a.ignore_index, a.scores, a.status, a.part, 0, 0, 1,
                                                                   &a.ctrl->cta_part[blockIdx.x], range_s);
    }
    // Which CTA's tiles straddle the start of this CTA's first chunk, and where that CTA's tiles begin: two numbers the
    // statistics behind the grid barrier need, found by a handful of 64-bit divisions -- done here by one thread of the
    // producer warp, which has issued its last tile and would only wait, instead of by every thread on the critical path behind
    // the barrier.  (Not in front of the scoring phase: the whole CTA would wait for it at the ring's initial barrier.)
    if (tid == kStepNT) {
        const bool wk = blockIdx.x != gridDim.x - 1 || gridDim.x == 1;
        const int nw = gridDim.x > 1 ? (int)gridDim.x - 1 : 1;
        const int q0 = a.nchunks / nw, big = nw - a.nchunks % nw;   // (as in phase B; this thread has time for the divisions)
        const int cf = wk ? (int)blockIdx.x * q0 + max(0, (int)blockIdx.x - big) : 0;
        const unsigned int G = gridDim.x, T = (unsigned int)a.ntiles;
        const unsigned int t0 = min((unsigned int)(cf * (kChunk / a.tile_rows)), T);
        unsigned int bstar = T > 0u ? t0 * G / T : 0u;   // (a first guess; the weights move it by a few CTAs at most)
        bstar = min(bstar, G - 1u);
        while (bstar + 1u < G && tile_begin(bstar + 1u, G, T) <= t0) ++bstar;
        while (bstar > 0u && tile_begin(bstar, G, T) > t0) --bstar;
        if (t0 >= T) bstar = G;   // (nothing in front of a chunk that starts behind the last tile -- cannot happen)
        b0_s[0] = bstar;
        b0_s[1] = bstar < G ? tile_begin(bstar, G, T) : T;
    }
    if (blockIdx.x == 0 && tid == 0) a.ctrl->times[1] = global_ns();
    if ((a.dbg_flags & 4) && tid == 0) {   // developer aid: when the LAST CTA had scored, and when each one had
        const unsigned long long now = global_ns();
        atomicMax(&a.ctrl->times[11], now);
        a.ctrl->k1_end[blockIdx.x] = (unsigned int)now;
    }
    grid.sync();
    if (warp == kStepNT / 32) return;  // the producer warp has nothing more to do (the barriers below are named, 256 threads)
    if (blockIdx.x == 0 && tid == 0) a.ctrl->times[2] = global_ns();

    const unsigned long long epoch = __ldcg(&a.ctrl->epoch);
    const bool f64 = F64 || a.cdf_mode == MERCURY_CDF_F64;   // (no walk: fp64 chunk sums, their scan, one CTA per draw)
    // ---- B: statistics (pytorch_collab.py:109-112), then: the last CTA is the walker and owns no chunks (with a single
    // CTA it is its own worker); the others each take a contiguous block of chunks
    const int nchunks = a.nchunks;
    const bool walker = blockIdx.x == gridDim.x - 1;
    const int nworkers = gridDim.x > 1 ? (int)gridDim.x - 1 : 1;
    const bool worker = !walker || gridDim.x == 1;
    const int q = a.chunks_q, r = a.chunks_r;   // (nchunks / nworkers and the remainder, from the host: no division behind the barrier)
    // (the r larger blocks at the END of the pool: the expensive chunks -- those in which the running sum changes binade,
    // crowded at the head of the pool -- then sit alone in their CTA)
    const int big0 = nworkers - r;
    const int c_first = worker ? (int)blockIdx.x * q + max(0, (int)blockIdx.x - big0) : 0;
    const int c_count = worker ? q + ((int)blockIdx.x >= big0 ? 1 : 0) : 0;
    const int per = a.tile_rows, npart = a.ntiles, parts_per_chunk = a.parts_per_chunk;
    float raw[8];
    auto load_raw = [&](int c) {
        const int64_t start = (int64_t)c * kChunk;
        const int len = (int)min((int64_t)kChunk, a.n - start);
        const int base = tid * 8;
        const float* wp = a.scores + start + base;
        if (base + 8 <= len && (reinterpret_cast<uintptr_t>(wp) & 15) == 0) {
            const float4 v0 = __ldcg(reinterpret_cast<const float4*>(wp)), v1 = __ldcg(reinterpret_cast<const float4*>(wp) + 1);
            raw[0] = v0.x; raw[1] = v0.y; raw[2] = v0.z; raw[3] = v0.w; raw[4] = v1.x; raw[5] = v1.y; raw[6] = v1.z; raw[7] = v1.w;
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) raw[k] = base + k < len ? __ldcg(wp + k) : 0.f;
        }
    };
    if (c_count > 0) load_raw(c_first);
    // the walker sums the head of the pool serially while the workers build their summaries (warp 1 below): its scores
    const int len0 = (int)min((int64_t)kChunk, a.n);
    const int H = min(kHeadMax, len0) & ~(kSub2 - 1);
    const float ema_prev = *a.ema_state;
    double tot = 0.0, pre = 0.0;
    double own_part[2] = {0.0, 0.0};   // this warp's chunk: its (at most 64) tile sums, two per lane, requested up front
    if (warp < min(c_count, 8)) {
        const int p0 = (c_first + warp) * parts_per_chunk;
        if (lane < parts_per_chunk && p0 + lane < npart) own_part[0] = __ldcg(a.part + p0 + lane);
        if (lane + 32 < parts_per_chunk && p0 + lane + 32 < npart) own_part[1] = __ldcg(a.part + p0 + lane + 32);
    }
    {
        // the CTA whose tiles straddle the start of this CTA's first chunk: the partial sums of the CTAs in front of it
        // count whole, its own tiles in front of the chunk one by one (warp 7)
        // (32-bit arithmetic: at most 2^16 tiles and 2^10 CTAs; 64-bit divisions are long instruction sequences)
        const unsigned int G = gridDim.x, T = (unsigned int)npart;
        const unsigned int t0 = min((unsigned int)(c_first * parts_per_chunk), T);
        const unsigned int bstar = b0_s[0];   // (found in front of the scoring phase, see there)
        double straddle = 0.0;
        if (warp == kStepNT / 32 - 1 && c_count > 0 && bstar < G)
            for (unsigned int i = b0_s[1] + lane; i < t0; i += 32) straddle += __ldcg(a.part + i);
        for (int b = tid; b < (int)G; b += kStepNT) {   // (fixed order: every CTA gets the same bits)
            const double v = __ldcg(&a.ctrl->cta_part[b]);
            tot += v;
            if ((unsigned int)b < bstar) pre += v;
        }
        if (warp == kStepNT / 32 - 1) pre += straddle;
        tot = warp_sum(tot); pre = warp_sum(pre);
        if (lane == 0) { s2.dred[0][warp] = tot; s2.dred[1][warp] = pre; }
        if (warp < min(c_count, 8)) {
            const double o = warp_sum(own_part[0] + own_part[1]);
            if (lane == 0) s2.own[warp] = o;
        }
        wsync();
        tot = pre = 0.0;
#pragma unroll
        for (int i = 0; i < kStepNT / 32; ++i) { tot += s2.dred[0][i]; pre += s2.dred[1][i]; }
    }
    double n_stat = (double)a.n;
    if (a.peers) {   // sharded pool: the statistics are those of the WHOLE pool -- CTA 0 exchanges the sums, everybody else takes its word
        __shared__ double gs[kMaxShards];
        if (blockIdx.x == 0) {
            PeerBox* const my_box = a.peers[a.rank];
            const unsigned int sum_tag = my_box->seq_sum + 1u;   // (the walker advances seq_sum once the step's sums are in)
            const int par = sum_tag & 1u;
            if (tid < a.nshards) {
                const unsigned long long bits = (unsigned long long)__double_as_longlong(tot);
                peer_store(&a.peers[tid]->sum_hi[par][a.rank], (bits >> 32 << 32) | sum_tag);
                peer_store(&a.peers[tid]->sum_lo[par][a.rank], (bits << 32) | sum_tag);
                unsigned long long hi, lo;
                const bool ok = peer_wait(&my_box->sum_hi[par][tid], sum_tag, hi, my_box) &
                                peer_wait(&my_box->sum_lo[par][tid], sum_tag, lo, my_box);
                if (!ok) atomicOr(a.status, MERCURY_ST_PEER_TIMEOUT);
                gs[tid] = __longlong_as_double((long long)((hi >> 32 << 32) | (lo >> 32)));
                a.sums_all[tid] = gs[tid];
            }
            wsync();
            tot = 0.0;
            for (int i = 0; i < a.nshards; ++i) tot += gs[i];   // (same order on every rank)
            if (tid == 0) {
                const unsigned long long bits = (unsigned long long)__double_as_longlong(tot);
                st_relaxed_u64(&a.ctrl->gsum[0], (1ull << 63) | (bits >> 32));
                st_relaxed_u64(&a.ctrl->gsum[1], (1ull << 63) | (bits & 0xffffffffull));
            }
        } else {
            if (tid == 0) {
                unsigned long long g0, g1;
                const unsigned long long t0 = global_ns();
                for (unsigned int it = 0;; ++it) {
                    g0 = ld_relaxed_u64(&a.ctrl->gsum[0]);
                    g1 = ld_relaxed_u64(&a.ctrl->gsum[1]);
                    if ((g0 & g1) >> 63) break;
                    if ((it & 63u) == 63u && global_ns() - t0 > kStepSpinNs) { atomicOr(a.status, MERCURY_ST_INTERNAL); break; }
                }
                gs[0] = __longlong_as_double((long long)(((g0 & 0xffffffffull) << 32) | (g1 & 0xffffffffull)));
            }
            wsync();
            tot = gs[0];
        }
        n_stat = (double)a.n_global;
    }
    const float mean = (float)(tot / n_stat);
    const float al = (float)a.ema_alpha, one_minus_a = (float)(1.0 - a.ema_alpha);
    const float ema = a.ema_first ? mean : __fadd_rn(__fmul_rn(al, ema_prev), __fmul_rn(one_minus_a, mean));
    const float shift = __fmul_rn((float)a.smooth_alpha, ema);
    const float total = (float)(tot + n_stat * (double)shift);
    if (blockIdx.x == 0 && tid == 0) {
        a.ctrl->times[14] = global_ns();
        a.stats[MERCURY_STAT_MEAN] = mean; a.stats[MERCURY_STAT_EMA] = ema;
        a.stats[MERCURY_STAT_SHIFT] = shift; a.stats[MERCURY_STAT_TOTAL] = total;
        a.stats[5] = a.stats[6] = a.stats[7] = 0.f;
    }
    for (int k = 0; k < c_count; ++k) {
        double own;
        if (k < 8) own = s2.own[k];
        else {  // (only with very few CTAs for the chunk count: not the common case)
            wsync();
            if (warp == 0) {
                const int p0 = (c_first + k) * parts_per_chunk;
                double o = 0.0;
                for (int i = p0 + lane; i < min(p0 + parts_per_chunk, npart); i += 32) o += __ldcg(a.part + i);
                o = warp_sum(o);
                if (lane == 0) s2.own[0] = o;
            }
            wsync();
            own = s2.own[0];
            wsync();
        }
        float cur[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) cur[i] = raw[i];
        if (k + 1 < c_count) load_raw(c_first + k + 1);   // next chunk's scores travel while this one is reduced
        if (f64) sum_f64_chunk(a, c_first + k, cur, shift, total, s2, epoch, k == 0);
        else {
            if (tid == 0 && !a.peers) a.approx[c_first + k] = pre;   // (for the draws' speculation, spec_chunk2: no ordering needed)
            summarize2_chunk(a, c_first + k, cur, pre, own, shift, total, s2, epoch, k == 0);
        }
        pre += own;
    }
    if (blockIdx.x == 0 && tid == 0) a.ctrl->times[15] = global_ns();
    if (blockIdx.x == 0 && tid == 0) a.ctrl->times[3] = global_ns();
    if (walker) {
        // ---- C: the head of the pool (warp 1: the reference's own loop over the first candidates, in which the running sum
        // changes binade about ten times), the wait for the workers' summaries (thread 0), then the walk
        if (warp == 1 && !f64) {
            uint32_t hb = 0u;
            if (H > 0) {
                // all loads first (8 x 16 bytes per lane in flight), then the arithmetic
                float4 hv[kHeadMax / 128];
                const bool vec = (reinterpret_cast<uintptr_t>(a.scores) & 15) == 0;
#pragma unroll
                for (int rr = 0; rr < kHeadMax / 128; ++rr) {
                    const int i = (rr * 32 + lane) * 4;
                    if (i < H) {
                        if (vec) hv[rr] = __ldcg(reinterpret_cast<const float4*>(a.scores + i));
                        else hv[rr] = make_float4(__ldcg(a.scores + i), __ldcg(a.scores + i + 1), __ldcg(a.scores + i + 2), __ldcg(a.scores + i + 3));
                    }
                }
                uint32_t* hp = reinterpret_cast<uint32_t*>(smem) + (kStepSmemBytes / 4 - kHeadMax);   // (the stage / event area: unused until the walk)
#pragma unroll
                for (int rr = 0; rr < kHeadMax / 128; ++rr) {
                    const int i = (rr * 32 + lane) * 4;
                    if (i < H)
                        *reinterpret_cast<uint4*>(hp + i) = make_uint4(f2b(__fdiv_rn(__fadd_rn(hv[rr].x, shift), total)), f2b(__fdiv_rn(__fadd_rn(hv[rr].y, shift), total)),
                                                                       f2b(__fdiv_rn(__fadd_rn(hv[rr].z, shift), total)), f2b(__fdiv_rn(__fadd_rn(hv[rr].w, shift), total)));
                }
                __syncwarp();
                const uint32_t haddr = smem_u32(hp);
                for (int blk = 0; blk < H / kSub2; ++blk) hb = serial32(hb, haddr + (uint32_t)blk * (kSub2 * 4u));
            }
            wst.env.head_k = H / kSub2; wst.env.head_bits = hb;   // (every lane stores the same words)
            if (lane == 0) a.ctrl->times[9] = global_ns();
        }
        wsync();
        if (f64) walker_f64_body(a, smem, wst, ema, epoch);
        else walker_body(a, smem, wst, shift, total, ema, epoch);
    }
    // ---- D: draws.  What a draw needs from the walker comes in its two words; the pool statistics are in registers
    // already (every CTA computed the same bits), so draw_resolve reads both from shared memory instead of global memory.
    float* cdf_s = reinterpret_cast<float*>(smem);
    __shared__ float dstats[MERCURY_STAT_COUNT];
    __shared__ float dtotal;
    __shared__ double drun0, dtot64;
    DrawArgs da = a.d;
    da.src.stats = dstats;
    da.total_ptr = &dtotal;
    // Speculation while the walk is under way (this CTA has nothing else to do): spec_chunk2 below.
    int chat = -1;
    if (!f64 && !a.peers && !walker && (int64_t)blockIdx.x < a.d.b && gridDim.x > 1) {
        if ((a.dbg_flags & 4) && blockIdx.x == 0 && tid == 0) a.ctrl->evt[97] = global_ns();
        chat = spec_chunk2(a.approx, nchunks, a.n, a.scores, a.d.u[blockIdx.x], shift, total, s2.dred[1], cdf_s);
        if ((a.dbg_flags & 4) && blockIdx.x == 0 && tid == 0) a.ctrl->evt[98] = global_ns();
    }
    for (int64_t j = blockIdx.x; j < a.d.b; j += gridDim.x) {
        wsync();   // (the walker's shared memory, and the previous draw's, are free again)
        if (tid == 0) {
            unsigned long long v, w;
            const unsigned long long t0 = global_ns();
            for (unsigned int it = 0;; ++it) {
                v = ld_relaxed_u64(&a.ctrl->dinfo[j]);
                w = ld_relaxed_u64(&a.ctrl->dinfo2[j]);
                if ((v & w) >> 63) {
                    if (!a.peers || ((v >> 62) & 1ull)) break;             // (not a sharded pool, or another rank's draw)
                    const unsigned long long x = ld_relaxed_u64(&a.ctrl->dinfo3[j]);
                    if (x >> 63) { dinfo3_s = x & ~(1ull << 63); break; }
                }
                if ((it & 63u) == 63u && global_ns() - t0 > kStepSpinNs) { atomicOr(a.status, MERCURY_ST_INTERNAL); v = 0ull; break; }
            }
            dinfo_s = v;
            dtotal = b2f((uint32_t)w);
            if (f64 && (v >> 63)) { __threadfence(); drun0 = __ldcg(&a.ctrl->dapprox[j]); dtot64 = __ldcg(&a.ctrl->dtotal[j]); }
            dstats[MERCURY_STAT_SHIFT] = shift; dstats[MERCURY_STAT_TOTAL] = total;
        }
        double uj = a.d.u[j];
        wsync();
        const unsigned long long v = dinfo_s;
        if (a.peers && (v >> 63)) uj = ((v >> 62) & 1ull) ? -1.0 : __longlong_as_double((long long)dinfo3_s);
        if ((a.dbg_flags & 4) && blockIdx.x == 0 && tid == 64) a.ctrl->evt[106] = global_ns();
        if (!(v >> 63)) return;  // timed out (reported): leave the outputs alone rather than write garbage
        if (f64) draw_resolve_f64<true>(da, j, uj, (int)((v >> 32) & 0x7fffffffu), drun0, dtot64, ds);
        else {
            const int chunk = (int)((v >> 32) & 0x7fffffffu);
            const bool hit = j == (int64_t)blockIdx.x && chunk == chat;
            if ((a.dbg_flags & 4) && tid == 0 && j == (int64_t)blockIdx.x) atomicAdd(reinterpret_cast<unsigned long long*>(&a.ctrl->evt[99]), hit ? 0x100000001ull : 1ull);
            draw_resolve<true>(da, j, uj, chunk, (uint32_t)v, true, ds, cdf_s, hit);
        }
    }
    if (blockIdx.x == 0 && tid == 0) a.ctrl->times[8] = global_ns();
}
