// CPU check of mercury_b200/csrc/chain.cuh (and of the draw kernel's in-chunk search, modelled below): the chunked "speculate, summarise, walk the spine,
// resolve the dirty chunks" scheme that cdf.cu runs on the GPU is emulated here sequentially and
// must reproduce the serial fp32 running sum (ATen's `sum += val`, reference pytorch_collab.py:114)
// bit for bit.  Build: g++ -O2 -ffp-contract=off -I mercury_b200/csrc tests/host/chain_host_test.cpp
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "chain.cuh"

using namespace mercury;

static uint32_t f2b(float f) { uint32_t b; memcpy(&b, &f, 4); return b; }
static float b2f(uint32_t b) { float f; memcpy(&f, &b, 4); return f; }

static std::vector<uint32_t> serial_chain(const std::vector<float>& p) {
    std::vector<uint32_t> out(p.size());
    volatile float s = 0.f;
    for (size_t i = 0; i < p.size(); ++i) { s += p[i]; out[i] = f2b(s); }
    return out;
}

// Sequential model of resolve_chunk(): phase loop over one chunk from an exact start state.
static uint32_t resolve_chunk(const float* p, int len, uint32_t bits, uint32_t* out) {
    int pos = 0;
    while (pos < len) {
        uint32_t E = phase_of(bits), S = bits - (E << 23);
        int cross = -1;
        // "parallel" part: inclusive composition of pairs from pos on, applied to S
        Pair acc = identity_pair();
        uint32_t prevS = S;
        for (int k = pos; k < len; ++k) {
            const Pair slow = addend_pair(f2b(p[k]), E), fast = addend_pair_fast(f2b(p[k]), E);
            if (slow.d0 != fast.d0 || slow.d1 != fast.d1) {
                printf("addend_pair_fast mismatch: p=%08x E=%u slow=(%u,%u) fast=(%u,%u)\n", f2b(p[k]), E, slow.d0, slow.d1, fast.d0, fast.d1);
                exit(2);
            }
            acc = compose(acc, fast);
            uint32_t Sk = apply(acc, S);
            if (Sk >= kPhaseLimit) { cross = k; break; }
            if (out) out[k] = bits_of(E, Sk);
            prevS = Sk;
        }
        if (cross < 0) return bits_of(E, prevS);
        float s = b2f(bits_of(E, prevS)) + p[cross];   // the one real fp32 add of this phase
        bits = f2b(s);
        if (out) out[cross] = bits;
        pos = cross + 1;
        if (pos == len) return bits;
    }
    return bits;
}

struct Summary { uint32_t e; Pair q[2]; };

static bool run_case(const char* name, const std::vector<float>& p, int chunk, double bias) {
    const size_t n = p.size();
    std::vector<uint32_t> ref = serial_chain(p);
    const size_t nchunks = (n + chunk - 1) / chunk;
    // 1. approximate prefix per chunk (fp64), as the stats kernel produces it
    std::vector<double> approx(nchunks + 1, 0.0);
    for (size_t c = 0; c < nchunks; ++c) {
        double s = 0;
        for (size_t i = c * chunk; i < n && i < (c + 1) * chunk; ++i) s += p[i];
        approx[c + 1] = approx[c] + s;
    }
    // 2. per-chunk summaries for the two candidate phases (parallel on the GPU)
    std::vector<Summary> sum(nchunks);
    for (size_t c = 0; c < nchunks; ++c) {
        float guess = (float)(approx[c] * bias);
        uint32_t e = phase_of(f2b(guess));
        sum[c].e = e;
        for (int v = 0; v < 2; ++v) {
            Pair acc = identity_pair();
            for (size_t i = c * chunk; i < n && i < (c + 1) * chunk; ++i) {
                const Pair slow = addend_pair(f2b(p[i]), e + v), fast = addend_pair_fast(f2b(p[i]), e + v);
                if (slow.d0 != fast.d0 || slow.d1 != fast.d1) {
                    printf("addend_pair_fast mismatch: p=%08x E=%u slow=(%u,%u) fast=(%u,%u)\n", f2b(p[i]), e + v, slow.d0, slow.d1, fast.d0, fast.d1);
                    exit(2);
                }
                acc = compose(acc, fast);
            }
            sum[c].q[v] = acc;
        }
    }
    // 3. spine walk
    std::vector<uint32_t> start(nchunks);
    std::vector<uint8_t> dirty(nchunks, 0);
    uint32_t bits = 0;
    size_t ndirty = 0;
    for (size_t c = 0; c < nchunks; ++c) {
        start[c] = bits;
        uint32_t E = phase_of(bits), S = bits - (E << 23);
        bool clean = false;
        if (E == sum[c].e || E == sum[c].e + 1) {
            uint32_t Sn = apply(sum[c].q[E - sum[c].e], S);
            if (Sn < kPhaseLimit) { bits = bits_of(E, Sn); clean = true; }
        }
        if (!clean) {
            dirty[c] = 1; ++ndirty;
            int len = (int)std::min<size_t>(chunk, n - c * chunk);
            bits = resolve_chunk(&p[c * chunk], len, bits, nullptr);
        }
    }
    // 4. materialise every chunk from its start state
    std::vector<uint32_t> got(n);
    for (size_t c = 0; c < nchunks; ++c) {
        int len = (int)std::min<size_t>(chunk, n - c * chunk);
        resolve_chunk(&p[c * chunk], len, start[c], &got[c * chunk]);
    }
    size_t bad = 0, first = 0;
    for (size_t i = 0; i < n; ++i) if (got[i] != ref[i]) { if (!bad) first = i; ++bad; }
    bool ok = bad == 0 && bits == ref[n - 1];
    printf("%-28s n=%-9zu chunk=%-5d dirty=%-4zu %s", name, n, chunk, ndirty, ok ? "OK\n" : "FAIL");
    if (!ok) printf(" bad=%zu first=%zu got=%08x ref=%08x end=%08x/%08x\n", bad, first, got[first], ref[first], bits, ref[n - 1]);
    return ok;
}

// Sequential model of draw_kernel's search inside one chunk (cdf.cu, "(b) the slot inside that chunk"): rounds of
// [one map per 64-candidate sub-chunk for the current phase -> scan from the sub-chunk where the exact sum is known
// -> either u falls before the phase change (walk its sub-chunk serially, done) or walk the sub-chunk with the
// change and repeat behind it].  Must return the number of candidates whose normalised running sum is < u.
static int draw_slot_model(const float* p, int len, uint32_t bits, float total, double u, int* rounds_out) {
    const int kSub = 64, kSubs = 32;
    std::vector<float> pv(kSub * kSubs, 0.f);
    for (int i = 0; i < len; ++i) pv[i] = p[i];
    int cur = 0, rounds = 0;
    for (;;) {
        ++rounds;
        const uint32_t E = phase_of(bits), S = bits - (E << 23);
        Pair subm[32];
        for (int k = 0; k < kSubs; ++k) {
            Pair m = identity_pair();
            for (int i = 0; i < kSub; ++i) m = compose(m, addend_pair_fast(f2b(pv[k * kSub + i]), E));
            subm[k] = m;
        }
        uint32_t s_out[32];
        Pair inc = identity_pair();
        for (int k = 0; k < kSubs; ++k) { if (k >= cur) inc = compose(inc, subm[k]); s_out[k] = apply(inc, S); }
        int f = kSubs;
        for (int k = cur; k < kSubs; ++k) if (s_out[k] >= kPhaseLimit) { f = k; break; }
        int target = -1; bool final_walk = false;
        for (int k = cur; k < f; ++k)
            if (!((double)(b2f(bits_of(E, s_out[k])) / total) < u)) { target = k; final_walk = true; break; }
        if (target < 0 && f < kSubs) target = f;
        if (target < 0) { *rounds_out = rounds; return len - 1; }
        float acc = b2f(bits_of(E, target > cur ? s_out[target - 1] : S));
        int cnt = 0;
        for (int i = 0; i < kSub; ++i) { acc = acc + pv[target * kSub + i]; cnt += ((double)(acc / total) < u) ? 1 : 0; }
        if (final_walk || cnt < kSub) { *rounds_out = rounds; return std::min(target * kSub + cnt, len - 1); }
        if (target + 1 >= kSubs) { *rounds_out = rounds; return len - 1; }
        cur = target + 1;
        bits = f2b(acc);
    }
}

static bool run_draw_case(const char* name, const std::vector<float>& p, std::mt19937_64& rng) {
    const int chunk = 2048;
    const size_t n = p.size();
    std::vector<uint32_t> ref = serial_chain(p);
    const float total = b2f(ref[n - 1]);
    if (!(total > 0.f) || std::isinf(total)) return true;
    const size_t nchunks = (n + chunk - 1) / chunk;
    std::uniform_real_distribution<double> U(0.0, 1.0);
    size_t bad = 0, tried = 0;
    int max_rounds = 0;
    for (int t = 0; t < 600; ++t) {
        // a uniform, and the chunk the device picks for it: the number of chunks whose last value is below u
        double u = t < 200 ? U(rng) * ((double)std::min<size_t>(n, 4096) / (double)n) : U(rng);
        if (t == 0) u = 0.0;
        if (t == 1) u = 1.0 - std::ldexp(1.0, -53);
        size_t c = 0;
        while (c < nchunks - 1 && (double)(b2f(ref[std::min(n, (c + 1) * chunk) - 1]) / total) < u) ++c;
        const int len = (int)std::min<size_t>(chunk, n - c * chunk);
        int rounds = 0;
        const int got = draw_slot_model(&p[c * chunk], len, c ? ref[c * chunk - 1] : 0u, total, u, &rounds);
        int want = 0;
        for (int i = 0; i < len; ++i) want += ((double)(b2f(ref[c * chunk + i]) / total) < u) ? 1 : 0;
        want = std::min(want, len - 1);
        max_rounds = std::max(max_rounds, rounds);
        ++tried;
        if (got != want) { if (!bad) printf("draw mismatch %s: u=%.17g chunk=%zu got=%d want=%d\n", name, u, c, got, want); ++bad; }
    }
    printf("%-28s n=%-9zu draws=%zu max rounds=%d %s\n", name, n, tried, max_rounds, bad ? "FAIL" : "OK");
    return bad == 0;
}

int main(int argc, char** argv) {
    size_t big = argc > 1 ? (size_t)atoll(argv[1]) : 1000000;
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::normal_distribution<double> G(0.0, 1.0);
    bool ok = true;
    for (size_t n : {(size_t)1, (size_t)7, (size_t)320, (size_t)5000, (size_t)50001, big}) {
        std::vector<float> p(n);
        double tot;
        // uniform-ish weights, normalised (the reference's p = w / sum(w))
        tot = 0; for (auto& v : p) { v = (float)(U(rng) * 3 + 0.01); tot += v; }
        for (auto& v : p) v = (float)(v / (float)tot);
        ok &= run_case("uniform normalised", p, 2048, 1.0 - 1e-4);
        ok &= run_case("uniform normalised c=256", p, 256, 1.0 - 1e-4);
        ok &= run_draw_case("draw: uniform normalised c=256", p, rng);
        // raw (un-normalised) heavy tail
        for (auto& v : p) v = (float)std::exp(G(rng) * 3.0);
        ok &= run_case("lognormal sigma=3 raw", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: lognormal sigma=3 raw", p, rng);
        // CE-like: loss + 0.5*mean, normalised
        tot = 0; for (auto& v : p) { v = (float)(-std::log(U(rng) + 1e-12) + 1.15); tot += v; }
        for (auto& v : p) v = (float)(v / (float)tot);
        ok &= run_case("CE-like normalised", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: CE-like normalised", p, rng);
        // constant weights: every add rounds the same way (worst systematic drift), many exact ties
        for (auto& v : p) v = 1.0f / (float)n;
        ok &= run_case("constant 1/n", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: constant 1/n", p, rng);
        for (auto& v : p) v = 0.75f;
        ok &= run_case("constant 0.75 (ties)", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: constant 0.75 (ties)", p, rng);
        // powers of two and half-ulps: tie storms
        for (size_t i = 0; i < n; ++i) p[i] = std::ldexp(1.0f, -(int)(rng() % 30));
        ok &= run_case("random powers of two", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: random powers of two", p, rng);
        // leading zeros, denormals, tiny then large
        for (size_t i = 0; i < n; ++i) {
            double r = U(rng);
            p[i] = i < n / 8 ? 0.f : (r < 0.2 ? (float)(1e-42 * r) : (r < 0.6 ? (float)(1e-30 * r) : (float)r));
        }
        ok &= run_case("zeros+denormals+mixed", p, 2048, 1.0 - 1e-4);
        ok &= run_draw_case("draw: zeros+denormals+mixed", p, rng);
        // bad speculation on purpose: the scheme must stay exact when every guess is wrong
        for (auto& v : p) v = (float)(U(rng));
        ok &= run_case("wrong guesses (bias 4x)", p, 2048, 4.0);
        ok &= run_case("wrong guesses (bias 0)", p, 2048, 0.0);
    }
    printf(ok ? "ALL OK\n" : "FAILURES\n");
    return ok ? 0 : 1;
}
