// Exact, order-independent arithmetic for the serial fp32 running sum that
// torch.multinomial's CPU kernel builds (`sum += val`, the reference's pytorch_collab.py:114).
//
// Idea (restated as a 2-state automaton).  Write the running sum by its IEEE
// bits.  For non-negative fp32 values, bits(s) = (E << 23) + S with "phase" E = max(expfield,1)-1
// and S in [0, 2^24): while the sum stays in one phase, its ulp is fixed (2^(E-149)) and
// `s = fl(s + p)` is the INTEGER update  S += a + b,  where p/ulp = a + r (a integer, r in [0,1)),
// b = [r > 1/2], and on an exact tie (r == 1/2) b = (S + a) & 1 (round half to even).  So each
// addend acts on S as  S -> S + d[S & 1]  for a pair d = (d0, d1) that depends only on p and E.
// Such maps compose associatively:  (A then B)[pi] = A[pi] + B[(pi + A[pi]) & 1]  -- any
// parallel scan order gives the serial answer.  A phase ends at the first addend that would
// push S to >= 2^24; that single add is done in real fp32 and defines the next phase.
//
// Everything here is __host__ __device__ so the same code is exercised by the CPU test
// (tests/host/chain_host_test.cpp) and by the kernels in cdf.cu.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MB_HD __host__ __device__ __forceinline__
#else
#define MB_HD inline
#endif

namespace mercury {

// Saturation cap for the integer state: anything >= 2^24 means "left the phase".
static constexpr uint32_t kPhaseLimit = 1u << 24;
static constexpr uint32_t kSatCap = 1u << 25;

struct Pair {  // the map S -> S + d[S & 1]
    uint32_t d0, d1;
};

MB_HD uint32_t sat(uint32_t x) { return x < kSatCap ? x : kSatCap; }

MB_HD Pair identity_pair() { return Pair{0u, 0u}; }

// (a then b)
MB_HD Pair compose(Pair a, Pair b) {
    Pair r;
    r.d0 = sat(a.d0 + ((a.d0 & 1u) ? b.d1 : b.d0));
    r.d1 = sat(a.d1 + (((a.d1 + 1u) & 1u) ? b.d1 : b.d0));
    return r;
}

MB_HD uint32_t apply(Pair a, uint32_t S) { return S + ((S & 1u) ? a.d1 : a.d0); }

// Phase of a non-negative finite fp32 given by its bits.
MB_HD uint32_t phase_of(uint32_t bits) {
    uint32_t ef = bits >> 23;
    return (ef > 0u ? ef : 1u) - 1u;
}

// The pair by which addend `pbits` (non-negative finite fp32 bits) acts in phase E.
MB_HD Pair addend_pair(uint32_t pbits, uint32_t E) {
    if (pbits == 0u) return Pair{0u, 0u};
    uint32_t ef = pbits >> 23;
    uint32_t M = pbits & 0x7fffffu;
    if (ef != 0u) M |= 0x800000u; else ef = 1u;       // denormal: no hidden bit, exponent of 2^-126
    int sh = (int)(E + 1u) - (int)ef;                  // log2(ulp_phase / ulp_addend)
    if (sh <= 0) {                                     // addend is a whole number of ulps
        int up = -sh;
        uint32_t a = up >= 2 ? kSatCap : sat(M << up);
        return Pair{a, a};
    }
    if (sh >= 25) return Pair{0u, 0u};                 // below half an ulp
    uint32_t a = M >> sh;
    uint32_t rem = M & ((1u << sh) - 1u);
    uint32_t half = 1u << (sh - 1);
    if (rem > half) return Pair{a + 1u, a + 1u};
    if (rem < half) return Pair{a, a};
    return Pair{a + (a & 1u), a + 1u - (a & 1u)};      // tie: result must be even
}

// The same pair through fp32 arithmetic (about a third of the instructions): p / ulp is p times
// a power of two, exact in fp32, so its floor and fraction are one FRND and one FADD away.
// Needs the scale 2^(149-E) to be a normal float, i.e. E >= 22; below that (running sums under
// 2^-104) the integer version above is used.
MB_HD float mb_bits2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
static constexpr uint32_t kFastPhaseMin = 22u;

#if defined(__CUDACC__)
static __device__ __noinline__ Pair addend_pair_outofline(uint32_t pbits, uint32_t E) { return addend_pair(pbits, E); }
#endif

MB_HD Pair addend_pair_fast(uint32_t pbits, uint32_t E) {
#if defined(__CUDA_ARCH__)
    if (E < kFastPhaseMin) return addend_pair_outofline(pbits, E);  // rare: keep it out of the hot loops
#else
    if (E < kFastPhaseMin) return addend_pair(pbits, E);
#endif
    const float scale = mb_bits2f((276u - E) << 23);   // 2^(149 - E)
    float t = mb_bits2f(pbits) * scale;                // p / ulp, exact (inf / tiny at the extremes)
    t = t < 33554432.f ? t : 33554432.f;               // cap at 2^25 (also catches inf and NaN)
    const float fl = floorf(t);
    const float fr = t - fl;                           // exact
    const uint32_t a = (uint32_t)fl;
    if (fr == 0.5f) return Pair{a + (a & 1u), a + 1u - (a & 1u)};  // tie: result must be even
    const uint32_t d = a + (fr > 0.5f ? 1u : 0u);
    return Pair{d, d};
}

// bits of the running sum from (phase, S)
MB_HD uint32_t bits_of(uint32_t E, uint32_t S) { return (E << 23) + S; }

}  // namespace mercury
