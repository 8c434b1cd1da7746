// Internal (non-ABI) entry points shared between the translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mercury_b200.h"

namespace mercury {

constexpr int kCdfChunk = 2048;  // candidates per CDF chunk (one CTA's worth)

// part_sums / per_out (optional): where the layout allows it, K1 also writes the sum of every `*per_out`
// consecutive scores (fp64, fixed order); *per_out == 0 means it did not.
int score_dispatch(const void* logits, int dtype, const int64_t* labels, int64_t n, int64_t c, int64_t row_stride,
                   int64_t ignore_index, float* scores, int32_t* status, cudaStream_t st, double* part_sums = nullptr,
                   int* per_out = nullptr, bool chained = false);  // chained: first kernel of the whole-step chain (PDL)

size_t cdf_ws_bytes(int64_t n);
int check_ws(void* ws, size_t have, size_t need, const char* who);

size_t update_ws_bytes(int64_t n);
double* update_ws_part_sums(void* ws, int64_t n);

// the whole step as one cooperative kernel; 1 = not applicable to this call (run the kernel chain instead)
int fused_step_launch(const mercury_update_args& a, void* ws, cudaStream_t st);

int pool_stats_launch(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds, float* ema_state, int ema_first,
                      double ema_alpha, double smooth_alpha, float* stats, void* ws, cudaStream_t st, bool update_layout);

// statistics + exact boundaries + draw over p_i = (scores_i + shift) / total
int update_tail_launch(const mercury_update_args& a, const float* scores, const double* part, int per, void* ws, cudaStream_t st);

}  // namespace mercury
