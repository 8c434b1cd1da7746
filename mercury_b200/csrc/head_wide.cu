// K1 with a WIDE classifier head folded in: scores[i] = CE(features[i] @ W^T + bias, labels[i]) without ever writing
// the logits.  Replaces `out = self.linear(out)` of the reference's Bottleneck ResNets (nn.Linear(512 * expansion,
// num_classes) = 2048 -> C, /root/reference/pytorch_model.py:39-64, 78, 94-96) followed by F.cross_entropy(...,
// reduction='none') (pytorch_collab.py:102).  This is the one dense contraction on the path, so it runs on the
// 5th-generation tensor cores:
//   * operands: features [n, d] and weight [c, d], both bf16 with d contiguous (K-major for the MMA), staged by TMA
//     (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 4-stage shared-memory ring, 128 x 64 and 256 x 64 per stage;
//   * tcgen05.mma (cta_group::1, kind::f16, M = 128, N = 256, K = 16) issued by ONE thread, fp32 accumulators in
//     tensor memory, double-buffered (2 x 256 columns): the MMAs of class tile j+1 run while tile j is drained;
//   * epilogue: four warps, one thread per row of the 128-row tile (TMEM lane = row): tcgen05.ld 32 columns at a time,
//     bias added, online log-sum-exp and the label's logit kept in registers across the class tiles; one score per row.
// A persistent grid (one CTA per SM) loops over the row tiles.  Roofline: tensor (2 * n * d * c flops).
#include <cuda.h>
#include <math.h>

#include <mutex>

#include "common.cuh"
#include "internal.cuh"

namespace mercury {
namespace {

// Two shapes of the same kernel: class tiles of 256 (4 stages; the 1000-class head, tensor-bound) and of 64 (8 stages; heads
// of up to a few dozen classes, where the kernel is a stream of features bound by HBM and a 256-wide tile would spend 25 x
// the useful MMA work and re-read a 32 KB weight tile per stage).
constexpr int kBM = 128, kBK = 64, kUmmaK = 16, kMaxStages = 8;
constexpr int kThreads = 192;                        // warp 0: TMA, warp 1: MMA + TMEM, warps 2..5: epilogue
constexpr uint32_t kABytes = kBM * kBK * 2;
constexpr int kMaxClasses = 4096;                    // bias staged in shared memory
constexpr float kLog2e = 1.4426950408889634f;

struct __align__(8) Bars {
    uint64_t full[kMaxStages], empty[kMaxStages], tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }

// bounded wait: a protocol mistake must end in a trap (an error the host sees), never in a GPU that stops answering
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    for (uint32_t it = 0;; ++it) {
        uint32_t done;
        asm volatile(
            "{\n.reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.b32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(a), "r"(parity) : "memory");
        if (done) return;
        if (it > (1u << 26)) __trap();
    }
}

__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* tm, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}

// shared-memory matrix descriptor, K-major, 128-byte swizzle: rows of 128 bytes, groups of 8 rows 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | (1ull << 16) /* LBO (unused with this swizzle) */ | (64ull << 32) /* SBO = 1024 B */ |
           (1ull << 46) /* descriptor version */ | (2ull << 61) /* SWIZZLE_128B */;
}
// instruction descriptor: D fp32, A and B bf16, both K-major, M = 128, N = bn
constexpr uint32_t idesc_for(int bn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
}

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {   // arrives on `bar` when every MMA issued so far has finished
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

template <int kBN, int kStages>
__global__ void __launch_bounds__(kThreads, 1)
head_wide_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, int64_t n, int d, int c,
                 const float* __restrict__ bias, const int64_t* __restrict__ labels, int64_t ignore_index,
                 float* __restrict__ scores, int32_t* status) {
    constexpr uint32_t kBBytes = kBN * kBK * 2, kStageBytes = kABytes + kBBytes;
    constexpr int kTmemCols = 2 * kBN;               // two accumulators (a power of two >= 32)
    constexpr uint32_t kIdesc = idesc_for(kBN);
    static_assert(kStages <= kMaxStages && (kTmemCols == 128 || kTmemCols == 512), "shapes");
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // (1024-byte alignment for the 128-byte swizzle; the launch asks for 1 KB of slack)
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* stage_a = smem;                                   // [kStages][128 x 64 bf16]
    unsigned char* stage_b = smem + (size_t)kStages * kABytes;       // [kStages][256 x 64 bf16]
    float* bias_s = reinterpret_cast<float*>(smem + (size_t)kStages * kStageBytes);   // [n_tiles * 256]
    const int num_n_tiles = (c + kBN - 1) / kBN, num_k = (d + kBK - 1) / kBK;
    Bars* bars = reinterpret_cast<Bars*>(bias_s + num_n_tiles * kBN);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t num_m_tiles = (n + kBM - 1) / kBM;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_b) : "memory");
        for (int s = 0; s < kStages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&bars->tmem_full[s], 1); mbar_init(&bars->tmem_empty[s], 128); }
        mbar_fence_init();
    }
    if (warp == 1) {   // tensor memory: all 512 columns, allocated (and later freed) by this warp
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)), "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = threadIdx.x; i < num_n_tiles * kBN; i += kThreads) bias_s[i] = (bias && i < c) ? bias[i] : 0.f;
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ===== TMA producer (one thread)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t mt = blockIdx.x; mt < num_m_tiles; mt += gridDim.x)
                for (int nt = 0; nt < num_n_tiles; ++nt)
                    for (int kb = 0; kb < num_k; ++kb) {
                        mbar_wait_bounded(&bars->empty[stage], phase ^ 1u);
                        mbar_arrive_expect_tx(&bars->full[stage], kStageBytes);
                        tma_load_2d(stage_a + (size_t)stage * kABytes, &tm_a, kb * kBK, (int)(mt * kBM), &bars->full[stage]);
                        tma_load_2d(stage_b + (size_t)stage * kBBytes, &tm_b, kb * kBK, nt * kBN, &bars->full[stage]);
                        if (++stage == kStages) { stage = 0; phase ^= 1u; }
                    }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread)
        if (lane == 0) {
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int64_t mt = blockIdx.x; mt < num_m_tiles; mt += gridDim.x)
                for (int nt = 0; nt < num_n_tiles; ++nt) {
                    mbar_wait_bounded(&bars->tmem_empty[acc], acc_phase ^ 1u);   // the epilogue has drained this accumulator
                    fence_after();
                    const uint32_t tmem_d = tmem_base + (uint32_t)acc * kBN;
                    for (int kb = 0; kb < num_k; ++kb) {
                        mbar_wait_bounded(&bars->full[stage], phase);
                        fence_after();
                        const uint64_t ad = umma_desc_sw128(smem_u32(stage_a + (size_t)stage * kABytes));
                        const uint64_t bd = umma_desc_sw128(smem_u32(stage_b + (size_t)stage * kBBytes));
#pragma unroll
                        for (int k = 0; k < kBK / kUmmaK; ++k)   // 16 bf16 = 32 bytes further along K: +2 in the address field
                            umma_f16(tmem_d, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), kIdesc, (kb | k) ? 1u : 0u);
                        umma_commit(&bars->empty[stage]);          // the stage is free once these MMAs have read it
                        if (++stage == kStages) { stage = 0; phase ^= 1u; }
                    }
                    umma_commit(&bars->tmem_full[acc]);              // the accumulator is complete
                    if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
                }
        }
    } else {
        // ===== epilogue: warp w may touch TMEM lanes 32 * (w % 4) .. + 31; one thread per row of the tile
        const int q = warp & 3;
        const int row_in_tile = q * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int64_t mt = blockIdx.x; mt < num_m_tiles; mt += gridDim.x) {
            const int64_t row = mt * kBM + row_in_tile;
            const int64_t y = row < n ? labels[row] : 0;
            const bool ok = row < n && y != ignore_index && y >= 0 && y < c;
            if (row < n && !ok && y != ignore_index) atomicOr(status, MERCURY_ST_BAD_LABEL);
            float m = -INFINITY, l = 0.f, xy = 0.f;   // running maximum, sum of 2^((x - m) log2 e), the label's logit
            for (int nt = 0; nt < num_n_tiles; ++nt) {
                mbar_wait_bounded(&bars->tmem_full[acc], acc_phase);
                fence_after();
#pragma unroll 1
                for (int ch = 0; ch < kBN / 32; ++ch) {
                    const int c0 = nt * kBN + ch * 32;
                    if (c0 >= c) break;                                  // (warp-uniform)
                    uint32_t v[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * kBN + ch * 32), v);
                    float x[32];
                    float cm = -INFINITY;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        x[i] = c0 + i < c ? __uint_as_float(v[i]) + bias_s[c0 + i] : -INFINITY;
                        cm = fmaxf(cm, x[i]);
                    }
                    if (y >= c0 && y < c0 + 32) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) if (c0 + i == y) xy = x[i];
                    }
                    const float mn = fmaxf(m, cm);
                    const float hi = mn * kLog2e;
                    float s0 = 0.f, s1 = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        s0 += ex2_approx(fmaf(x[i], kLog2e, -hi));
                        s1 += ex2_approx(fmaf(x[i + 1], kLog2e, -hi));
                    }
                    l = fmaf(l, ex2_approx((m - mn) * kLog2e), s0 + s1);   // (m == -inf: the factor is 0 and l is 0)
                    m = mn;
                }
                fence_before();
                mbar_arrive(&bars->tmem_empty[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            }
            if (row < n) scores[row] = ok ? (logf(l) + m) - xy : 0.f;
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
}

// cuTensorMapEncodeTiled through the runtime (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// [rows, d] bf16, row stride in elements, box = 64 columns x box_rows rows, 128-byte swizzle, zero fill outside
int make_map(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t d, int64_t row_stride, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) { set_last_error("mercury_score_head_wide: cuTensorMapEncodeTiled is not available"); return MERCURY_ERR_CUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)d, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_stride * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_last_error("mercury_score_head_wide: cuTensorMapEncodeTiled failed (%d)", (int)r); return MERCURY_ERR_CUDA; }
    return MERCURY_OK;
}

}  // namespace
}  // namespace mercury

using namespace mercury;

extern "C" int mercury_score_head_wide(const void* features, int64_t n, int64_t d, int64_t row_stride, const void* weight,
                                       int64_t weight_stride, const float* bias, int64_t c, const int64_t* labels,
                                       int64_t ignore_index, float* scores, int32_t* status, void* stream) {
    MB_CHECK_ARG(n >= 0 && d >= 8 && c >= 1 && c <= kMaxClasses, "mercury_score_head_wide: need n >= 0, d >= 8, 1 <= c <= %d", kMaxClasses);
    if (n == 0) return MERCURY_OK;
    MB_CHECK_ARG(features && weight && labels && scores && status, "mercury_score_head_wide: null pointer");
    MB_CHECK_ARG(row_stride >= d && weight_stride >= d && (row_stride % 8) == 0 && (weight_stride % 8) == 0,
                 "mercury_score_head_wide: row strides must be >= d and multiples of 8 elements (16 bytes)");
    MB_CHECK_ARG(((reinterpret_cast<uintptr_t>(features) | reinterpret_cast<uintptr_t>(weight)) & 15) == 0,
                 "mercury_score_head_wide: features and weight must be 16-byte aligned");
    MB_CHECK_ARG(n < ((int64_t)1 << 31) && d < ((int64_t)1 << 24), "mercury_score_head_wide: n or d too large for the tile coordinates");
    const bool narrow = c <= 128;          // (up to two class tiles of 64)
    const int bn = narrow ? 64 : 256, stages = narrow ? 8 : 4;
    CUtensorMap tm_a, tm_b;
    int rc = make_map(&tm_a, features, n, d, row_stride, kBM);
    if (!rc) rc = make_map(&tm_b, weight, c, d, weight_stride, bn);
    if (rc) return rc;
    const int num_n_tiles = (int)((c + bn - 1) / bn);
    const size_t smem = 1024 + (size_t)stages * (kABytes + (size_t)bn * kBK * 2) + (size_t)num_n_tiles * bn * 4 + sizeof(Bars) + 64;
    auto kern = narrow ? head_wide_kernel<64, 8> : head_wide_kernel<256, 4>;
    rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem, "head_wide_kernel smem attribute");
    if (rc) return rc;
    const int sms = device_sm_count();
    const int64_t tiles = (n + kBM - 1) / kBM;
    const int grid = (int)min_i64(tiles, sms > 0 ? sms : kSMs);
    kern<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(tm_a, tm_b, n, (int)d, (int)c, bias, labels, ignore_index, scores, status);
    MB_CHECK_LAUNCH("head_wide_kernel");
    return MERCURY_OK;
}
