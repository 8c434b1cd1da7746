// ABI housekeeping: version, per-thread error text, and the whole-step entry point that chains
// K1 -> stats -> exact CDF boundaries -> fused draw/gather (reference pytorch_collab.py:102-117).
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"
#include "internal.cuh"

namespace mercury {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("MERCURY_PDL"); return !(e && e[0] == '0'); }();
    return on;
}

namespace {
constexpr int kMaxDevicesTracked = 64;
std::mutex g_attr_mu;
std::map<std::pair<int, const void*>, size_t> g_attr_done;  // (device ordinal, function) -> bytes already granted
int g_sm_count[kMaxDevicesTracked] = {};
}  // namespace

int ensure_dynamic_smem(const void* func, size_t bytes, const char* what) {
    if (bytes <= 48 * 1024) return MERCURY_OK;  // within the default limit: nothing to opt in to
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, what);
    std::lock_guard<std::mutex> lock(g_attr_mu);
    size_t& have = g_attr_done[std::make_pair(dev, func)];
    if (have >= bytes) return MERCURY_OK;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return cuda_fail(e, what);
    have = bytes;
    return MERCURY_OK;
}

int device_sm_count() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevicesTracked) return 0;
    std::lock_guard<std::mutex> lock(g_attr_mu);
    if (g_sm_count[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        g_sm_count[dev] = v;
    }
    return g_sm_count[dev];
}

int env_knob(const char* name, int dflt, int lo, int hi) {
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    char* end = nullptr;
    const long v = strtol(e, &end, 10);
    if (end == e || *end != '\0' || v < lo || v > hi) {
        fprintf(stderr, "mercury_b200: ignoring %s=%s (expected an integer in [%d, %d])\n", name, e, lo, hi);
        return dflt;
    }
    return (int)v;
}

int cuda_fail(cudaError_t e, const char* what) {
    set_last_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return MERCURY_ERR_CUDA;
}

}  // namespace mercury

using namespace mercury;

extern "C" int mercury_abi_version(void) { return MERCURY_ABI_VERSION; }
extern "C" const char* mercury_last_error(void) { return g_err; }

extern "C" size_t mercury_update_samples_workspace_bytes(int64_t n, int64_t b) {
    (void)b;
    return update_ws_bytes(n);
}

extern "C" int mercury_update_samples(const mercury_update_args* a, void* ws, size_t ws_bytes, void* stream) {
    MB_CHECK_ARG(a != nullptr, "mercury_update_samples: null argument block");
    MB_CHECK_ARG(a->n >= 1, "mercury_update_samples: empty pool");
    MB_CHECK_ARG(a->n <= ((int64_t)1 << 24), "number of categories cannot exceed 2^24");
    MB_CHECK_ARG(a->b >= 0, "mercury_update_samples: negative draw count");
    MB_CHECK_ARG(a->stats && a->ema_state && a->status, "mercury_update_samples: stats / ema_state / status are required");
    MB_CHECK_ARG(a->b == 0 || (a->uniforms && a->idx && a->nprob && a->seed),
                 "mercury_update_samples: uniforms / idx / nprob / seed are required when b > 0");
    MB_CHECK_ARG(!a->global_sums || (a->n_global_sums >= 1 && a->n_global >= a->n && a->n_bounds == 0),
                 "mercury_update_samples: sharded call needs n_global_sums >= 1, n_global >= n and no pool-size list");
    MB_CHECK_ARG(!a->global_sums || a->n <= ((int64_t)8192 * 2048), "mercury_update_samples: shard too large");
    MB_CHECK_ARG(!a->draw_only || a->scores_in, "mercury_update_samples: draw_only needs scores_in");
    MB_CHECK_ARG(a->cdf_mode == MERCURY_CDF_EXACT_SERIAL_F32 || a->cdf_mode == MERCURY_CDF_F64, "mercury_update_samples: unknown cdf_mode %d",
                 a->cdf_mode);
    MB_CHECK_ARG(a->cdf_mode == MERCURY_CDF_EXACT_SERIAL_F32 || (!a->global_sums && !a->peer_boxes),
                 "mercury_update_samples: MERCURY_CDF_F64 excludes global_sums and peer_boxes");
    if (a->peer_boxes) {
        MB_CHECK_ARG(!a->global_sums && !a->draw_only && a->n_bounds == 0,
                     "mercury_update_samples: peer_boxes excludes global_sums, draw_only and a pool-size list");
        MB_CHECK_ARG(a->n_shards >= 1 && a->n_shards <= 64 && a->shard_rank >= 0 && a->shard_rank < a->n_shards,
                     "mercury_update_samples: 1..64 shards, rank inside");
        MB_CHECK_ARG(a->sums_all && a->totals_all && (a->b == 0 || a->local_uniforms) && a->n_global >= a->n,
                     "mercury_update_samples: sharded step needs sums_all, totals_all, local_uniforms and n_global >= n");
        MB_CHECK_ARG(a->n <= ((int64_t)8192 * 2048), "mercury_update_samples: shard too large");
    }
    const float* scores = a->scores_in;
    if (!scores) {
        MB_CHECK_ARG(a->logits && a->labels && a->scores, "mercury_update_samples: logits / labels / scores are required");
        MB_CHECK_ARG(a->c >= 1 && a->row_stride >= a->c, "mercury_update_samples: bad class count or row stride");
        MB_CHECK_ARG(a->dtype == MERCURY_F32 || a->dtype == MERCURY_BF16, "mercury_update_samples: bad dtype");
    }
    MB_CHECK_ARG(a->n_bounds >= 0 && a->n_bounds <= 64, "mercury_update_samples: at most 64 pool sizes");
    for (int i = 0; i < a->n_bounds; ++i) {
        MB_CHECK_ARG(a->h_bounds[i] >= 1 && a->h_bounds[i] <= a->n && (i == 0 || a->h_bounds[i] > a->h_bounds[i - 1]),
                     "mercury_update_samples: pool sizes must be ascending in [1, n]");
    }
    MB_CHECK_ARG(a->n_bounds == 0 || a->h_bounds[a->n_bounds - 1] == a->n, "mercury_update_samples: last pool size must equal n");
    int rc = check_ws(ws, ws_bytes, update_ws_bytes(a->n), "mercury_update_samples");
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (!scores) {  // small rows: the whole step is one persistent kernel
        MB_CHECK_ARG(a->logits && a->labels && a->scores, "mercury_update_samples: logits / labels / scores are required");
        MB_CHECK_ARG(a->c >= 1 && a->row_stride >= a->c, "mercury_update_samples: bad class count or row stride");
        MB_CHECK_ARG(a->dtype == MERCURY_F32 || a->dtype == MERCURY_BF16, "mercury_update_samples: bad dtype");
        rc = fused_step_launch(*a, ws, st);
        if (rc <= 0) return rc;
    }
    const double* part = nullptr;
    int per = 0;
    if (!scores) {
        double* part_buf = update_ws_part_sums(ws, a->n);
        rc = score_dispatch(a->logits, a->dtype, a->labels, a->n, a->c, a->row_stride, a->ignore_index, a->scores, a->status, st,
                            part_buf, &per, true);
        if (rc) return rc;
        scores = a->scores;
        part = part_buf;
    }
    return update_tail_launch(*a, scores, part, per, ws, st);
}
