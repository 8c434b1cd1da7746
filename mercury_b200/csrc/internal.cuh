// Internal (non-ABI) entry points shared between the translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mercury_b200.h"

namespace mercury {

constexpr int kCdfChunk = 2048;  // candidates per CDF chunk (one CTA's worth)

int score_dispatch(const void* logits, int dtype, const int64_t* labels, int64_t n, int64_t c, int64_t row_stride,
                   int64_t ignore_index, float* scores, int32_t* status, cudaStream_t st);

size_t cdf_ws_bytes(int64_t n);
int check_ws(void* ws, size_t have, size_t need, const char* who);

int pool_stats_launch(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds, float* ema_state, int ema_first,
                      double ema_alpha, double smooth_alpha, float* stats, void* ws, cudaStream_t st);

// summarize + spine + draw over p_i = (scores_i + shift) / total; `ws` already holds the
// approximate chunk prefixes written by pool_stats_launch
int update_draw_launch(const mercury_update_args& a, const float* scores, void* ws, cudaStream_t st);

}  // namespace mercury
