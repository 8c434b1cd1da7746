// K3 -- gather the drawn candidates into the training batch, with the importance weights
// N*p_i and the backward seed of the reweighted mean loss.
// Replaces reference pytorch_collab.py:116-117 (fancy-index gathers, `* N`) and the gradient that
// :137 (torch.div), :145 (torch.mean) and :148 (backward) deliver to the per-sample losses.
#include "common.cuh"
#include "internal.cuh"

namespace mercury {
namespace {

constexpr int kGatherNT = 128;

__global__ void __launch_bounds__(kGatherNT) gather_kernel(const float* __restrict__ p, const int64_t* __restrict__ idx,
                                                           int64_t b, int64_t n, const unsigned char* __restrict__ data,
                                                           int64_t row_bytes, const int64_t* __restrict__ labels,
                                                           const int64_t* __restrict__ index, unsigned char* __restrict__ out_data,
                                                           int64_t* __restrict__ out_labels, int64_t* __restrict__ out_index,
                                                           float* __restrict__ nprob, float* __restrict__ seed) {
    const int64_t j = blockIdx.x;
    int64_t pick = idx[j];
    pick = pick < 0 ? 0 : (pick >= n ? n - 1 : pick);
    if (threadIdx.x == 0) {
        const float np_ = __fmul_rn(p[pick], (float)n);                 // :116  losses_pros[idx] * N
        nprob[j] = np_;
        seed[j] = __fdiv_rn(__fdiv_rn(1.f, (float)b), np_);             // :145 mean -> 1/B ; :137 div -> / probs
        if (labels && out_labels) out_labels[j] = labels[pick];
        if (index && out_index) out_index[j] = index[pick];
    }
    if (data && out_data) {
        const unsigned char* srow = data + pick * row_bytes;
        unsigned char* drow = out_data + j * row_bytes;
        if (((reinterpret_cast<uintptr_t>(srow) | reinterpret_cast<uintptr_t>(drow) | (uintptr_t)row_bytes) & 15) == 0) {
            const uint4* s4 = reinterpret_cast<const uint4*>(srow);
            uint4* d4 = reinterpret_cast<uint4*>(drow);
            for (int64_t i = threadIdx.x; i < (row_bytes >> 4); i += kGatherNT) d4[i] = ldg_stream(s4 + i);
        } else {
            for (int64_t i = threadIdx.x; i < row_bytes; i += kGatherNT) drow[i] = srow[i];
        }
    }
}

// sharded pools: `idx` are GLOBAL pool positions; this rank owns [shard_lo, shard_lo + shard_n).
// Owned draws are gathered from the local shard, the others are zero-filled, so that a SUM
// all-reduce over the ranks assembles the batch without any host round trip.
__global__ void __launch_bounds__(kGatherNT) gather_owned_kernel(const int64_t* __restrict__ idx, int64_t shard_lo, int64_t shard_n,
                                                                 const unsigned char* __restrict__ data, int64_t row_bytes,
                                                                 const int64_t* __restrict__ labels, const int64_t* __restrict__ index,
                                                                 unsigned char* __restrict__ out_data, int64_t* __restrict__ out_labels,
                                                                 int64_t* __restrict__ out_index, int32_t* __restrict__ owned) {
    const int64_t j = blockIdx.x;
    const int64_t local = idx[j] - shard_lo;
    const bool mine = local >= 0 && local < shard_n;
    if (threadIdx.x == 0) {
        if (owned) owned[j] = mine ? 1 : 0;
        if (labels && out_labels) out_labels[j] = mine ? labels[local] : 0;
        if (index && out_index) out_index[j] = mine ? index[local] : 0;
    }
    if (data && out_data) {
        const unsigned char* srow = data + (mine ? local : 0) * row_bytes;
        unsigned char* drow = out_data + j * row_bytes;
        if (((reinterpret_cast<uintptr_t>(srow) | reinterpret_cast<uintptr_t>(drow) | (uintptr_t)row_bytes) & 15) == 0) {
            const uint4* s4 = reinterpret_cast<const uint4*>(srow);
            uint4* d4 = reinterpret_cast<uint4*>(drow);
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            for (int64_t i = threadIdx.x; i < (row_bytes >> 4); i += kGatherNT) d4[i] = mine ? ldg_stream(s4 + i) : z;
        } else {
            for (int64_t i = threadIdx.x; i < row_bytes; i += kGatherNT) drow[i] = mine ? srow[i] : (unsigned char)0;
        }
    }
}

// mean_j(loss_j / (N p_j)) = sum_j loss_j * seed_j: one CTA, pairwise tree in a fixed order (same bits run to run)
__global__ void __launch_bounds__(256) reweighted_mean_kernel(const float* __restrict__ losses, const float* __restrict__ seed, int64_t b,
                                                              float* __restrict__ out) {
    __shared__ float red[256];
    float acc = 0.f;
    for (int64_t j = threadIdx.x; j < b; j += 256) acc = __fmaf_rn(losses[j], seed[j], acc);
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] = __fadd_rn(red[threadIdx.x], red[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = red[0];
}

}  // namespace
}  // namespace mercury

using namespace mercury;

extern "C" int mercury_reweighted_mean(const float* losses, const float* seed, int64_t b, float* out, void* stream) {
    MB_CHECK_ARG(b >= 0 && out && (b == 0 || (losses && seed)), "mercury_reweighted_mean: null pointer or negative b");
    reweighted_mean_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(losses, seed, b, out);
    MB_CHECK_LAUNCH("reweighted_mean_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_gather_owned(const int64_t* idx, int64_t b, int64_t shard_lo, int64_t shard_n, const void* data,
                                    int64_t row_bytes, const int64_t* labels, const int64_t* index, void* out_data,
                                    int64_t* out_labels, int64_t* out_index, int32_t* owned, void* stream) {
    MB_CHECK_ARG(b >= 0 && shard_n >= 0 && shard_lo >= 0, "mercury_gather_owned: bad sizes");
    if (b == 0) return MERCURY_OK;
    MB_CHECK_ARG(idx != nullptr, "mercury_gather_owned: null idx");
    MB_CHECK_ARG(!data || row_bytes > 0, "mercury_gather_owned: row_bytes must be positive when data is given");
    gather_owned_kernel<<<(unsigned)b, kGatherNT, 0, static_cast<cudaStream_t>(stream)>>>(
        idx, shard_lo, shard_n, static_cast<const unsigned char*>(data), row_bytes, labels, index,
        static_cast<unsigned char*>(out_data), out_labels, out_index, owned);
    MB_CHECK_LAUNCH("gather_owned_kernel");
    return MERCURY_OK;
}

extern "C" int mercury_gather_reweight(const float* p, const int64_t* idx, int64_t b, int64_t n, const void* data,
                                       int64_t row_bytes, const int64_t* labels, const int64_t* index, void* out_data,
                                       int64_t* out_labels, int64_t* out_index, float* nprob, float* seed, void* stream) {
    MB_CHECK_ARG(b >= 0 && n >= 1, "mercury_gather_reweight: need b >= 0, n >= 1");
    if (b == 0) return MERCURY_OK;
    MB_CHECK_ARG(p && idx && nprob && seed, "mercury_gather_reweight: null pointer");
    MB_CHECK_ARG(!data || row_bytes > 0, "mercury_gather_reweight: row_bytes must be positive when data is given");
    gather_kernel<<<(unsigned)b, kGatherNT, 0, static_cast<cudaStream_t>(stream)>>>(
        p, idx, b, n, static_cast<const unsigned char*>(data), row_bytes, labels, index, static_cast<unsigned char*>(out_data),
        out_labels, out_index, nprob, seed);
    MB_CHECK_LAUNCH("gather_kernel");
    return MERCURY_OK;
}
