/* mercury_b200 -- C ABI of the B200-native importance-sampling hot path.
 *
 * Drop-in boundary.  The reference (AIoT-MLSys-Lab/Mercury) has no FFI of its own: its hot path
 * is the Python method Trainer.update_samples (pytorch_collab.py:89-117) plus three lines of
 * Trainer.train (:137,:145,:148) and the alternative Groupwise_Sampler (util.py:94-160), and all
 * arithmetic is delegated to PyTorch calls.  Each entry point below replaces one of those call
 * sites; the citation after "replaces" names it.  mercury_b200/ops.py binds them with ctypes and
 * INTEGRATION.md shows the stub a maintainer of the reference would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name starts with h_;
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*), performs no
 *    allocation, no host synchronisation and keeps no global state -- with these exceptions, each
 *    stated again at its declaration:
 *      * mercury_peer_create / _open / _close / _destroy are set-up calls: they allocate (cudaMalloc),
 *        clear and synchronise the device once; the per-step calls that use the boxes do not;
 *      * mercury_cdf_static works in a static device array shared by the process (one per device) and
 *        caches that array's address per device under a mutex on first use;
 *      * per device and kernel, the first launch of a kernel that needs more than 48 KB of dynamic shared
 *        memory (or a cooperative launch) sets the function attribute / queries occupancy once and
 *        remembers it in a mutex-protected table (host memory, no device allocation);
 *      * process-wide knobs read ONCE from the environment at first use and then fixed: MERCURY_PDL
 *        (0 = no programmatic dependent launch), MERCURY_K1_TILE / MERCURY_K1_CTAS (tile rows / grid of
 *        the small-row scoring kernel; validated, bad values are ignored), MERCURY_FUSED (-1 never,
 *        0 automatic, 1 always: the one-kernel step), MERCURY_STEP_PDL, MERCURY_STEP_CTAS, MERCURY_STEP_DBG,
 *        MERCURY_K1_WARP_CTAS, MERCURY_K1_WIDE_CTAS, MERCURY_K1_POLY and
 *        MERCURY_SPINE_DEBUG (development), MERCURY_PEER_WAIT_MS (bound of one peer wait);
 *      * mercury_test_set_guess_bias / mercury_test_set_fused / mercury_debug_step_times are TEST HOOKS:
 *        process-global atomics read at launch time, there so the parity tests can force the rare paths
 *        (mis-guessed binades, the fallback chain); a production caller never calls them;
 *  - scratch memory is supplied by the caller (`ws`, sized by the matching *_workspace_bytes);
 *    workspaces must be 256-byte aligned; a workspace may be reused by the next call on the
 *    same stream without clearing it;
 *  - return value: 0 on success, a negative MERCURY_ERR_* otherwise (mercury_last_error() gives
 *    the text for the calling thread);
 *  - data errors that torch reports by raising (bad label, negative / non-finite probability,
 *    zero total) are reported through a device status word (bit set MERCURY_ST_*), which the
 *    caller may read whenever it next synchronises.
 */
#ifndef MERCURY_B200_H
#define MERCURY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MERCURY_ABI_VERSION 3

enum { MERCURY_OK = 0, MERCURY_ERR_ARG = -1, MERCURY_ERR_CUDA = -2, MERCURY_ERR_WORKSPACE = -3 };

enum { MERCURY_F32 = 0, MERCURY_BF16 = 1 };

/* how the cumulative distribution is accumulated */
enum {
    MERCURY_CDF_EXACT_SERIAL_F32 = 0, /* bit-identical to ATen's serial `sum += val` in fp32 */
    MERCURY_CDF_F64 = 1               /* deterministic fp64 reduce-then-scan (not bit-identical) */
};

/* device status word bits */
enum {
    MERCURY_ST_BAD_LABEL = 1,  /* F.cross_entropy: "Target out of bounds"                          */
    MERCURY_ST_NEGATIVE = 2,   /* multinomial: "encountering probability entry < 0" (NaN included) */
    MERCURY_ST_NONFINITE = 4,  /* multinomial: "probability entry = infinity or NaN"               */
    MERCURY_ST_ZERO_SUM = 8,   /* multinomial: "sum of probabilities <= 0"                         */
    MERCURY_ST_TOPK_TIES = 16, /* exponential race: more exact ties at the k-th key than buffered  */
    MERCURY_ST_PEER_TIMEOUT = 32, /* peer exchange: a peer's word did not arrive within the bound   */
    MERCURY_ST_INTERNAL = 64,  /* a wait inside the one-kernel step ran into its 2 s bound (a bug)    */
    MERCURY_ST_GROUP_COUNT = 128 /* group-masked draw: the group has not the member count the host gave */
};

/* layout of the 8-float statistics block written by mercury_pool_stats */
enum {
    MERCURY_STAT_MEAN = 0,     /* mean(scores)              pytorch_collab.py:109 */
    MERCURY_STAT_EMA = 1,      /* EMAverage.value afterwards util.py:207-213      */
    MERCURY_STAT_SHIFT = 2,    /* alpha * ema                pytorch_collab.py:111 */
    MERCURY_STAT_TOTAL = 3,    /* sum(scores + shift)        pytorch_collab.py:112 */
    MERCURY_STAT_CHAIN_TOTAL = 4, /* last value of the serial fp32 running sum over p (whole-step call only) */
    MERCURY_STAT_COUNT = 8
};

int mercury_abi_version(void);
const char* mercury_last_error(void);

/* ---- K1: per-candidate importance score -----------------------------------------------------
 * replaces F.cross_entropy(output, label, reduction='none') at pytorch_collab.py:102 and
 * util.py:129.  logits [n, c] with `row_stride` elements between rows, dtype MERCURY_F32 or
 * MERCURY_BF16 (upcast in registers); labels int64 [n]; label == ignore_index scores 0; a label
 * outside [0, c) sets MERCURY_ST_BAD_LABEL and scores 0.  scores fp32 [n].  The logits are read
 * exactly once. */
int mercury_score(const void* logits, int dtype, const int64_t* labels, int64_t n, int64_t c,
                  int64_t row_stride, int64_t ignore_index, float* scores, int32_t* status,
                  void* stream);

/* ---- pool statistics, smoothing shift, normaliser --------------------------------------------
 * replaces torch.mean (:109), EMAverage.update (:110, util.py:207-213), `+ alpha*ema` (:111) and
 * torch.sum (:112).  `ema_state` is one device float holding EMAverage.value; `ema_first` != 0
 * means "first update copies".  With n_bounds > 0, `h_bounds` (HOST int64, ascending, last == n)
 * lists growing pool sizes and the EMA is updated once per prefix mean, reproducing the
 * reference's 10 updates per call (:95-110); with n_bounds == 0 there is one update on the
 * whole pool.  Writes stats[MERCURY_STAT_COUNT] (see enum) and updates *ema_state. */
size_t mercury_pool_stats_workspace_bytes(int64_t n);
int mercury_pool_stats(const float* scores, int64_t n, const int64_t* h_bounds, int n_bounds,
                       float* ema_state, int ema_first, double ema_alpha, double smooth_alpha,
                       float* stats, void* ws, size_t ws_bytes, void* stream);

/* p_i = (score_i + shift) / total, the vector the reference hands to torch.multinomial
 * (pytorch_collab.py:111-112), materialised. */
int mercury_normalize(const float* scores, int64_t n, const float* stats, float* p, void* stream);

/* ---- K2: cumulative distribution --------------------------------------------------------------
 * replaces the first half of torch.multinomial(p, B, replacement=True) at pytorch_collab.py:114
 * (ATen multinomial_with_replacement: cum_dist).  weights fp32 [n] (>= 0, any scale);
 * cdf fp32 [n] = running sum / total; *total = the running sum's last value.  n <= 2^24. */
size_t mercury_cdf_workspace_bytes(int64_t n);
int mercury_cdf(const float* weights, int64_t n, int mode, float* cdf, float* total,
                int32_t* status, void* ws, size_t ws_bytes, void* stream);

/* No counterpart in the reference.  mercury_cdf for hosts that cannot allocate device memory
 * (a plugin inside a runtime that owns the whole GPU, a small C/Go/Java binding with no
 * allocator of its own): same inputs, outputs and result as mercury_cdf, but the scratch is a
 * 208 KiB array built into the library (held on every GPU that loads it), so nothing is
 * allocated or freed, ever.  n <= MERCURY_STATIC_MAX_N, larger pools are refused with
 * MERCURY_ERR_ARG.  The first call on each device looks the array's address up; make that call
 * outside stream capture.  All calls on a device share the array: calls that overlap on the GPU
 * race, so keep them on one stream.  Nothing in the array needs clearing between calls. */
#define MERCURY_STATIC_MAX_N (1 << 22)
int mercury_cdf_static(const float* weights, int64_t n, int mode, float* cdf, float* total,
                       int32_t* status, void* stream);

/* replaces the second half of torch.multinomial(..., replacement=True) (:114): for each fp64
 * uniform u_j the first slot with (double)cdf[slot] >= u_j.  idx int64 [b]. */
int mercury_sample_with_replacement(const float* cdf, int64_t n, const double* uniforms, int64_t b,
                                    int64_t* idx, void* stream);

/* ---- K2': exponential race (Gumbel top-k), draws without replacement --------------------------
 * replaces torch.multinomial(p, k, replacement=False) and every n_sample == 1 draw (util.py:150):
 * keys_i = weights_i / float(-log1p(-u_i)), idx = the k largest keys in descending order.
 * uniforms fp64 [n].  k <= 2048. */
size_t mercury_sample_topk_workspace_bytes(int64_t n, int64_t k);
int mercury_sample_topk(const float* weights, int64_t n, const double* uniforms, int64_t k,
                        int64_t* idx, int32_t* status, void* ws, size_t ws_bytes, void* stream);

/* ---- group-masked draws ------------------------------------------------------------------------
 * replaces Groupwise_Sampler.__iter__'s body (util.py:140-157): mask = (tags == group); imp = importance[mask];
 * imp + np.mean(imp); / np.sum; torch.multinomial(torch.Tensor(imp), 1, replacement) and the nonzero() that maps the
 * winner back -- without compacting anything.  importance fp32 [n], tags int32 [n]; `count` = members of the group
 * (the host knows it: it wrote the tags; a mismatch sets MERCURY_ST_GROUP_COUNT).  p = (imp + mean) / sum in float64,
 * rounded once to fp32; member m (in position order) races with uniform m of its draw: uniforms fp64 [ndraws * count].
 * k == 1: ndraws independent single draws (idx int64 [ndraws], dataset positions); k > 1 (ndraws == 1): the k largest
 * keys in descending order (draw without replacement).  p_out fp32 [n] (optional): p at the members, 0 elsewhere. */
size_t mercury_groupwise_draw_workspace_bytes(int64_t n, int64_t ndraws);
int mercury_groupwise_draw(const float* importance, const int32_t* tags, int32_t group, int64_t n, int64_t count,
                           const double* uniforms, int64_t ndraws, int64_t k, int64_t* idx, float* p_out,
                           int32_t* status, void* ws, size_t ws_bytes, void* stream);

/* ---- K3: gather the drawn candidates + importance weights + backward seed ----------------------
 * replaces losses_pros[idx] * N, cat(datas)[idx], cat(labels)[idx], cat(index)[idx]
 * (pytorch_collab.py:116-117) and the gradient that torch.div / torch.mean / backward
 * (:137,:145,:148) send to the per-sample losses: seed_j = (1/b) / nprob_j.
 * p fp32 [n]; data [n, row_bytes] bytes (may be NULL, then out_data is ignored); labels / index
 * int64 [n] (may be NULL).  nprob, seed fp32 [b]. */
int mercury_gather_reweight(const float* p, const int64_t* idx, int64_t b, int64_t n,
                            const void* data, int64_t row_bytes, const int64_t* labels,
                            const int64_t* index, void* out_data, int64_t* out_labels,
                            int64_t* out_index, float* nprob, float* seed, void* stream);

/* torch.mean(torch.div(losses, probs)) of pytorch_collab.py:137,145 as sum_j losses_j * seed_j (seed from the call above):
 * losses, seed fp32 [b] -> out fp32 [1], summed in a fixed order. */
int mercury_reweighted_mean(const float* losses, const float* seed, int64_t b, float* out, void* stream);

/* Sharded pools (one process per GPU, pytorch_collab.py:282-287 gives each worker its own
 * presample shard): `idx` are GLOBAL pool positions, this rank owns [shard_lo, shard_lo+shard_n).
 * Rows of owned draws are gathered from the local shard, all other output rows are zero-filled
 * (labels / index 0, owned[j] = 0), so a SUM all-reduce across ranks assembles the batch that
 * pytorch_collab.py:116-117 gathers from a single pool.  owned int32 [b] may be NULL. */
int mercury_gather_owned(const int64_t* idx, int64_t b, int64_t shard_lo, int64_t shard_n,
                         const void* data, int64_t row_bytes, const int64_t* labels,
                         const int64_t* index, void* out_data, int64_t* out_labels,
                         int64_t* out_index, int32_t* owned, void* stream);

/* Test hook, no counterpart in the reference: scales every phase guess of the exact CDF (1.0 =
 * normal).  Guesses only choose which maps are precomputed, so any value must leave results
 * unchanged; the tests use it to force the mis-guess paths.  Process-global; not for production. */
void mercury_test_set_guess_bias(double bias);

/* Test hook, no counterpart in the reference: mercury_update_samples runs pools of small rows (<= 256 B of logits per
 * candidate, <= 2 Mi candidates, <= 1024 draws, one EMA update) as ONE cooperative kernel, everything else as a chain
 * of four kernels; both give the same bits.  mode 0 forces the chain, 1 the single kernel where it applies, -1 (default)
 * follows the environment (MERCURY_FUSED=0 turns the single kernel off).  Process-global; not for production.
 * mercury_debug_step_times copies to h_out128 (HOST, 128 values) what the last single-kernel step left in workspace
 * `ws`: [0..8] %globaltimer stamps in ns -- start, CTA 0 scored, grid barrier passed, CTA 0 published its summaries,
 * last CTA arrived (walker starts), walker loaded, walk done, draws published, CTA 0 done; [16..] clock64 << 4 | kind
 * after each step of the walk (MERCURY_STEP_DBG=1).  It synchronises the device: a developer aid. */
void mercury_test_set_fused(int mode);
int mercury_debug_step_times(const void* ws, int64_t n, unsigned long long* h_out128);

/* ---- sharded pools, totals-only exchange (north_star: "per-shard score totals and CDFs are exchanged
 * with NCCL allgather ... the global draw is resolved locally") ------------------------------------
 * No counterpart in the reference (its workers never exchange importance, pytorch_collab.py:282-287).
 * Per rank: mercury_score -> mercury_sum_scores -> all-gather of the G sums -> mercury_update_samples
 * with b = 0 draws on the local shard (global_sums/n_global set: whole-pool mean/EMA/shift/total; the
 * shard's own exact chain, whose last value lands in stats[MERCURY_STAT_CHAIN_TOTAL]) -> all-gather of
 * the G chain totals -> mercury_shard_select maps every uniform to (shard, uniform inside the shard)
 * -> mercury_update_samples again for the draws (uniforms < 0 = "not mine": all-zero outputs, so a SUM
 * all-reduce assembles the batch).  The two-level draw is exact for its own definition (oracle:
 * oracle.sharded_draw_two_level); it is not index-identical to torch.multinomial on the flat pool. */
int mercury_sum_scores(const float* scores, int64_t n, double* out_sum, void* ws, size_t ws_bytes, void* stream);
int mercury_shard_select(const float* shard_totals, int nshards, int rank, const double* uniforms, int64_t b,
                         double* local_uniforms, void* stream);

/* ---- the same exchange over peer memory instead of a collective (no counterpart in the reference) ----
 * Set-up, once per rank: mercury_peer_create allocates this rank's zeroed exchange box
 * (MERCURY_PEER_BOX_BYTES of device memory) and returns a 64-byte CUDA IPC handle for it; the host
 * passes the handles round (any channel), maps every peer's box with mercury_peer_open and keeps the G
 * box addresses, in rank order with its own box at [rank], in a DEVICE array `peer_boxes`.
 * mercury_peer_close unmaps a peer's box, mercury_peer_destroy frees the own one.
 * Per step: mercury_sum_scores_exchange = mercury_sum_scores + all-gather (sums_all fp64 [nshards] is
 * filled for every rank); mercury_shard_select_exchange = all-gather of the chain totals (my_total: device
 * fp32, e.g. &stats[MERCURY_STAT_CHAIN_TOTAL]; totals_all fp32 [nshards] out) + mercury_shard_select.
 * The kernels store into the peers' boxes and wait for the peers' stores: every rank must make the same
 * sequence of exchange calls, each on a GPU of its own, and reach each call within the bound of one wait
 * (4 s, or MERCURY_PEER_WAIT_MS read by mercury_peer_create, or mercury_peer_set_wait) -- a loop in which one
 * rank does minutes of extra work between steps (the reference's rank-0 evaluation, pytorch_collab.py:181)
 * needs a barrier in front of the next step, a larger bound, or the collective exchange instead.
 * A wait that runs out sets MERCURY_ST_PEER_TIMEOUT, marks the box dead (later waits give up at once) and
 * POISONS the step: local_uniforms = NaN, and the draws of that rank get idx -1 and NaN nprob / seed, which
 * survive a SUM all-reduce.  The host must treat the status as fatal for the step; after all ranks have
 * synchronised (barrier), mercury_peer_reset on every rank's own box (then another barrier) returns the
 * boxes to their initial state. */
#define MERCURY_PEER_BOX_BYTES 4096
#define MERCURY_PEER_HANDLE_BYTES 64
int mercury_peer_create(void** box, unsigned char* handle);
int mercury_peer_open(const unsigned char* handle, void** peer_box);
int mercury_peer_close(void* peer_box);
int mercury_peer_destroy(void* box);
int mercury_peer_set_wait(void* box, int64_t wait_ms, void* stream);   /* own box; 0 < wait_ms <= 3 600 000 */
int mercury_peer_reset(void* box, void* stream);                       /* own box; collective, see above     */
int mercury_sum_scores_exchange(const float* scores, int64_t n, void* const* peer_boxes, int nshards, int rank,
                                double* sums_all, int32_t* status, void* ws, size_t ws_bytes, void* stream);
int mercury_shard_select_exchange(const float* my_total, void* const* peer_boxes, int nshards, int rank,
                                  const double* uniforms, int64_t b, float* totals_all, double* local_uniforms,
                                  int32_t* status, void* stream);

/* ---- K1 with the classifier head folded in ---------------------------------------
 * replaces  out = self.linear(features)            pytorch_model.py:94-96 (ResNet: 512 -> num_classes)
 *           F.cross_entropy(out, label, reduction='none')              pytorch_collab.py:102
 * scores[i] = CE(features[i] @ weight^T + bias, labels[i]); the logits are never written.  features
 * [n, d] fp32 or bf16 (dtype as in mercury_score), row_stride in elements, rows 16-byte aligned, d a
 * multiple of 128; weight fp32 [c, d] row-major (nn.Linear.weight), bias fp32 [c] or NULL; 1 <= c <= 16
 * and c * d <= 51200 (the matrix lives in shared memory): the reference's 512 -> 10 head.  Accumulation is
 * fp32 in a fixed order that is not cuBLAS's: scores agree with the reference's to ~1e-6 absolute per unit
 * of |logit| (tests: rtol 2e-5, atol 2e-5).  Labels as in mercury_score. */
int mercury_score_head(const void* features, int dtype, int64_t n, int64_t d, int64_t row_stride,
                       const float* weight, const float* bias, int64_t c, const int64_t* labels,
                       int64_t ignore_index, float* scores, int32_t* status, void* stream);

/* The same for WIDE heads, on the tensor cores (the reference's Bottleneck ResNets end in
 * nn.Linear(512 * expansion, num_classes) = 2048 -> C, pytorch_model.py:39-64, 78, 94-96): features [n, d] bf16 and
 * weight [c, d] bf16 (nn.Linear.weight cast once), both with d contiguous, row strides in elements (multiples of 8),
 * 16-byte aligned; bias fp32 [c] or NULL; 1 <= c <= 4096.  TMA-staged operands, tcgen05.mma with fp32 accumulators in
 * tensor memory, online log-sum-exp in the epilogue: the [n, c] logits are never written.  Products are bf16 x bf16 in
 * fp32: scores agree with F.cross_entropy(F.linear(features.float(), weight.float(), bias), labels) to rtol/atol 2e-5.
 * Labels as in mercury_score. */
int mercury_score_head_wide(const void* features, int64_t n, int64_t d, int64_t row_stride, const void* weight,
                            int64_t weight_stride, const float* bias, int64_t c, const int64_t* labels,
                            int64_t ignore_index, float* scores, int32_t* status, void* stream);

/* ---- the whole step ---------------------------------------------------------------------------
 * replaces Trainer.update_samples' arithmetic (pytorch_collab.py:102-117) on an assembled pool:
 * score -> stats -> exact CDF -> b draws from caller-supplied uniforms -> gather + weights + seed.
 * Never materialises p or the CDF: only the chunks a draw lands in are rebuilt.  If `scores_in`
 * is non-NULL the scoring step is skipped and those scores are used (logits may then be NULL);
 * the sharded mode uses this after the all-gather.  Optional outputs may be NULL. */
typedef struct mercury_update_args {
    const void* logits; int dtype; const int64_t* labels; int64_t n, c, row_stride, ignore_index;
    const float* scores_in;           /* optional: precomputed scores [n]                        */
    const int64_t* h_bounds; int n_bounds;  /* see mercury_pool_stats                             */
    float* ema_state; int ema_first; double ema_alpha, smooth_alpha;
    const double* uniforms; int64_t b;      /* fp64 uniforms [b]                                  */
    const void* data; int64_t row_bytes; const int64_t* index;   /* payload [n,row_bytes], ids [n] */
    float* scores;                    /* out fp32 [n] (required unless scores_in given)          */
    float* stats;                     /* out fp32 [MERCURY_STAT_COUNT]                           */
    int64_t* idx;                     /* out int64 [b] pool positions                             */
    float* nprob; float* seed;        /* out fp32 [b]                                            */
    void* out_data; int64_t* out_labels; int64_t* out_index;    /* out, gathered rows             */
    int32_t* status;                  /* in/out device status word                               */
    /* sharded pools (optional, see mercury_shard_select): this call covers ONE shard of n candidates;   */
    const double* global_sums; int n_global_sums; /* per-shard sums of scores of the whole pool [n_global_sums] */
    int64_t n_global;                 /* candidates in the whole pool (statistics and N*p use it)    */
    int64_t index_base;               /* added to idx[]: global position of the shard's first candidate */
    int draw_only;                    /* != 0: only draw + gather, from the boundaries / statistics that the  */
                                      /* previous call with the same workspace, scores and stats left behind  */
    /* the whole sharded step in ONE call, exchanges over peer memory (optional; see mercury_peer_create):       */
    /* the same four kernels as on one GPU, the exchanges inside them: score -> summaries (CTA 0 stores this     */
    /* shard's sum of scores into every peer's box, all CTAs wait for the G sums) -> spine (stores the shard's   */
    /* chain total into every peer's box) -> draw (waits for the G totals, picks each uniform's shard, resolves  */
    /* those of this shard; others: all-zero outputs).  global_sums must be NULL, draw_only 0; n_global /        */
    /* index_base as above.  Ranks must call in step (see mercury_peer_create).                                  */
    void* const* peer_boxes;          /* DEVICE array of the n_shards exchange boxes, own box at [shard_rank]  */
    int n_shards, shard_rank;
    double* sums_all;                 /* out fp64 [n_shards]: every shard's sum of scores                      */
    float* totals_all;                /* out fp32 [n_shards]: every shard's chain total                        */
    double* local_uniforms;           /* out fp64 [b]: position inside this shard's CDF, or -1 (not mine)      */
    /* how the cumulative distribution is accumulated (ABI 3): MERCURY_CDF_EXACT_SERIAL_F32 (0, the default: indices     */
    /* bit-identical to torch.multinomial on the CPU for the same p and uniforms) or MERCURY_CDF_F64: cdf_i = float(fp64  */
    /* running total / fp64 total) -- deterministic and without the serial walk, a draw from the same distribution but   */
    /* not the same indices.  MERCURY_CDF_F64 excludes global_sums and peer_boxes.                                       */
    int cdf_mode;
} mercury_update_args;

size_t mercury_update_samples_workspace_bytes(int64_t n, int64_t b);
int mercury_update_samples(const mercury_update_args* args, void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MERCURY_B200_H */
