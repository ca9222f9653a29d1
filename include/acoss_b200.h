/*
 * acoss_b200.h -- C ABI of libacoss_b200.so: the B200 (sm_100a) replacement for the
 * native + numpy arithmetic of acoss's all-pairs Serra09 scoring path.
 *
 * Every entry point is plain C (pointers + sizes, no torch / C++ types), returns an
 * int status (0 = ok, <0 = error, message through acoss_last_error()) and never throws.
 * All pointers are HOST pointers unless stated; the library owns its device memory.
 * One context per (device, host thread); calls on one context are serialised by the caller.
 *
 * What each entry point replaces in the reference (paths relative to the acoss tree):
 *
 *   acoss_upload_songs      the per-process feature cache of Serra09.load_features
 *                           (benchmarking/Serra09.py:96-156: 'chroma' (bins, L) and
 *                           'gchroma' (bins,) per song), packed into one arena
 *   acoss_serra09_pairs     the per-pair body of Serra09.similarity
 *                           (benchmarking/Serra09.py:165-175): get_oti
 *                           (CRPUtils.py:109-136) -> np.roll (:167) -> get_csm
 *                           (CRPUtils.py:67-84) -> sliding_csm (:24-45) ->
 *                           csm_to_binary_mutual (:201-219) -> qmax / dmax
 *                           (pySeqAlign.pyx:14-23 -> SequenceAlignment.c:113-180)
 *   acoss_align_dense       the tail of EarlySNF.similarity (benchmarking/EarlySNF.py:86-90)
 *                           and EarlySNF_Old.py:200-218: a caller-supplied dense M x N
 *                           dissimilarity -> csm_to_binary[_mutual] -> qmax/dmax/swconstrained
 *   acoss_qmax / acoss_dmax / acoss_swconstrained
 *                           the Cython FFI itself (pySeqAlign.pyx:7-23): S uint8 [M*N],
 *                           same argument order and return value; the caller's float D
 *                           table is not needed (and not filled)
 *   acoss_snf_pairs         EarlySNF.similarity's fusion step (EarlySNF.py:44-57, 85-90;
 *                           SimilarityFusion.py:50-177, 202-277) + its binarise / qmax / dmax tail
 *   acoss_dp_packed         batched qmax/dmax/sw over caller-supplied bit-packed CRPs
 *   acoss_aggregate_songs   Serra09.load_features' block aggregation + global_chroma
 *                           (benchmarking/Serra09.py:24-28, 96-114)
 *   acoss_eval_ranks        the sort + rank loop of CoverAlgorithm.getEvalStatistics
 *                           (benchmarking/CoverAlgorithm.py:345-382)
 *   acoss_bind_score_matrices
 *                           the `self.Ds[type][i, j] = score` side effect of Serra09.similarity and the
 *                           `Ds += Ds.T` of all_pairwise (benchmarking/Serra09.py:174-175, 181-184;
 *                           CoverAlgorithm.py:180-182), kept in device memory for acoss_eval_ranks
 *
 * Without a counterpart in the reference: acoss_dp_packed_static (see below), the measurement
 * aids acoss_last_stage_ms / acoss_last_eval_ms / acoss_last_aggregate_stats / acoss_kernel_launches and the
 * profiling aids acoss_debug_*.
 */
#ifndef ACOSS_B200_H
#define ACOSS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct acoss_ctx acoss_ctx;

/* status codes */
#define ACOSS_OK 0
#define ACOSS_ERR_ARG (-1)     /* bad argument (null pointer, negative size, index out of range) */
#define ACOSS_ERR_CUDA (-2)    /* a CUDA runtime call or kernel failed */
#define ACOSS_ERR_STATE (-3)   /* songs not uploaded yet / context destroyed */
#define ACOSS_ERR_NOMEM (-4)

/* feature dtype of the arena handed to acoss_upload_songs */
#define ACOSS_F32 0            /* hpcp / mfcc_htk (preprocess/features.py:209-275,418-470) */
#define ACOSS_F64 1            /* crema (preprocess/features.py:277-296) */

/* flags for the scoring entry points */
#define ACOSS_FLAG_NO_OTI       (1u << 0) /* skip OTI + roll (mfcc path, Serra09.py:177-180) */
#define ACOSS_FLAG_DMAX_CLEAN   (1u << 1) /* dmax on a zeroed table (ChenFusion.py:66); default is the
                                             table left behind by qmax, as Serra09.py:173-175 does */
#define ACOSS_FLAG_NO_SWAP      (1u << 2) /* keep rows = first song of the pair (debug CRP dumps are then
                                             in the reference's orientation); by default the shorter song
                                             is put on the columns (the DPs are transpose-invariant) */
#define ACOSS_FLAG_NON_MUTUAL   (1u << 3) /* csm_to_binary instead of csm_to_binary_mutual (row kNN only) */
#define ACOSS_FLAG_WANT_SW      (1u << 4) /* also run the constrained Smith-Waterman DP */
#define ACOSS_FLAG_DEVICE_INPUT (1u << 8) /* acoss_eval_ranks: D is a device pointer on the context's device
                                             (e.g. a torch tensor's data_ptr()); everything else stays host */

const char* acoss_last_error(void);
int acoss_version(void);

/* Create / destroy a context bound to CUDA device `device`. Fails (never falls back to the
 * CPU) when there is no usable device. */
int acoss_create(acoss_ctx** out, int device);
int acoss_destroy(acoss_ctx* ctx);

/*
 * Upload all songs once. frames: frame-major arena [total_frames][bins] of `dtype`
 * (song s occupies rows offsets[s] .. offsets[s]+lengths[s]-1; this is chroma.T of the
 * reference's (bins, L) array, so no transposition happens on the device).
 * gchroma: [n_songs][bins] float64 (Serra09.py:24-28,101-102) or NULL when every call
 * will pass ACOSS_FLAG_NO_OTI.  bins <= 16.
 * `frames` and `gchroma` may be HOST or DEVICE pointers (e.g. the arena a multi-GPU run has just
 * all-gathered over NCCL: the copy is then device-to-device, no host bounce); the caller makes sure
 * the producer of a device buffer has finished.  offsets and lengths are always host arrays.
 */
int acoss_upload_songs(acoss_ctx* ctx, const void* frames, int dtype,
                       const int64_t* offsets, const int32_t* lengths,
                       int32_t n_songs, int32_t bins, const double* gchroma);

/*
 * Score P song pairs. pairs: [P][2] int32 song indices (i, j) exactly as
 * CoverAlgorithm.similarity receives them (CoverAlgorithm.py:117, Serra09.py:158-162).
 * m: embedding window (Serra09.py:87), kappa: neighbour fraction (0 < kappa < 1).
 * Outputs (each [P], any may be NULL):
 *   qmax_raw, dmax_raw, sw_raw : the DP maxima exactly as qmax()/dmax()/swconstrained()
 *                                return them (NOT yet divided by M+N, Serra09.py:174-175)
 *   oti_out                    : the optimal transposition index used
 *   shape_out [P][2]           : (M, N) = (L_i-m+1, L_j-m+1)
 * Debug CRP dump (crp_out may be NULL): pair p's packed CRP is written at word offset
 * crp_offsets[p]: rows x ceil(cols/32) uint32, bit t of word w of row r = CRP[r][32w+t].
 * With ACOSS_FLAG_NO_SWAP rows = M, cols = N; otherwise crp_swapped_out[p] (may be NULL)
 * tells whether the stored matrix is the transpose.
 */
int acoss_serra09_pairs(acoss_ctx* ctx, const int32_t* pairs, int64_t P,
                        int32_t m, double kappa, uint32_t flags,
                        float* qmax_raw, float* dmax_raw, float* sw_raw,
                        int32_t* oti_out, int32_t* shape_out,
                        uint32_t* crp_out, const int64_t* crp_offsets, int32_t* crp_swapped_out);

/*
 * Dense entry (EarlySNF.py:86-90): B dense dissimilarity matrices of `dtype` (ACOSS_F32 /
 * ACOSS_F64; the reference's fused matrix is float64), matrix b is rows[b] x cols[b]
 * row-major at csm + csm_offsets[b] (elements).  Binarises with
 * k_row = round(kappa*cols), k_col = round(kappa*rows) (CRPUtils.py:186-199,219) and
 * runs the DPs.  out: [B][3] = qmax, dmax, sw maxima (sw only with ACOSS_FLAG_WANT_SW, else 0).
 */
int acoss_align_dense(acoss_ctx* ctx, const void* csm, int dtype, const int64_t* csm_offsets,
                      const int32_t* rows, const int32_t* cols, int32_t B,
                      double kappa, uint32_t flags, float* out,
                      uint32_t* crp_out, const int64_t* crp_offsets);

/*
 * Similarity-network-fusion front end of EarlySNF (benchmarking/EarlySNF.py:44-57, 85-90 with
 * benchmarking/SimilarityFusion.py:50-177, 202-277: get_WCSMSSM, get_P, get_S, snf_ws(niters, reg_diag=True)),
 * followed by the EarlySNF tail (csm_to_binary_mutual -> qmax / dmax on one table).  Two feature kinds are fused:
 * the arena of acoss_upload_songs (kind 0, rolled by the OTI of the pair unless ACOSS_FLAG_NO_OTI) and a second
 * arena of the SAME songs (same offsets and lengths, e.g. the 13-bin mfcc sequence) uploaded with
 * acoss_upload_aux_songs; the reference's second kind, the scattering-transform SSMs, needs kymatio and is out of
 * scope.  K = int(kappa (M + N)) neighbours, float64 throughout.  Limits: M + N <= 2048, K <= 256.
 * Outputs per pair: raw qmax / dmax maxima, oti, shape (M, N); optionally minus-the-fused cross block
 * (fused_out + fused_offsets[p], M x N float64, row-major: the matrix the reference binarises) and its packed CRP.
 */
int acoss_upload_aux_songs(acoss_ctx* ctx, const void* frames, int dtype, int32_t bins);
int acoss_snf_pairs(acoss_ctx* ctx, const int32_t* pairs, int64_t P, int32_t m, double kappa, int32_t niters,
                    uint32_t flags, float* qmax_raw, float* dmax_raw, int32_t* oti_out, int32_t* shape_out,
                    double* fused_out, const int64_t* fused_offsets, uint32_t* crp_out, const int64_t* crp_offsets);

/*
 * Batched DPs over caller-supplied packed CRPs (same packing as the debug dump).
 * out: [B][3] = qmax, dmax, sw.
 */
int acoss_dp_packed(acoss_ctx* ctx, const uint32_t* crp, const int64_t* crp_offsets,
                    const int32_t* rows, const int32_t* cols, int32_t B,
                    uint32_t flags, float* out);

/*
 * acoss_dp_packed for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * The entry point calls no cudaMalloc / cudaFree / allocating library routine and copies nothing between host and
 * device: its scratch (plot records, item list, column-block hand-over records) is one file-scope __device__ array
 * of 4 202 496 bytes (csrc/noalloc.cu) that every GPU loading the library holds until the process exits.
 *   crp_dev   DEVICE pointer, packed plots (bit t of word w of row r = S[r][32 w + t]); plot b starts at word
 *             crp_offsets[b] and has rows[b] x ceil(cols[b] / 32) words
 *   crp_offsets, rows, cols   HOST arrays of B entries (read before the call returns)
 *   out_dev   DEVICE pointer to 3 * B floats: [0, B) qmax, [B, 2B) dmax, [2B, 3B) sw (0 without
 *             ACOSS_FLAG_WANT_SW); same values as acoss_dp_packed gives for the same plots and flags
 *   stream    a cudaStream_t (NULL = the default stream); the call only enqueues two kernels on it and does not
 *             synchronise: the caller waits on the stream before reading out_dev
 * Limits (ACOSS_ERR_ARG beyond them): B <= 64 plots per call, rows and cols <= 16384, and at most 65536 rows in
 * total over the plots wider than 128 columns.  Every call on a device shares the array, which keeps the previous
 * call's contents (what must be fresh is rewritten by the call itself): calls that overlap on the GPU race, so
 * serialise them.  The first call on a device also looks up the array's device address (cudaGetSymbolAddress);
 * make that call outside a graph capture.
 */
int acoss_dp_packed_static(int32_t device, const uint32_t* crp_dev, const int64_t* crp_offsets,
                           const int32_t* rows, const int32_t* cols, int32_t B,
                           uint32_t flags, float* out_dev, void* stream);

/*
 * Drop-in equivalents of pySeqAlign.qmax / dmax / swconstrained (pySeqAlign.pyx:7-23):
 * S is the unpacked uint8 matrix, row-major.  qmax/dmax: S is M x N (rows, cols);
 * swconstrained: S is N x M (its first int is the row count, as in
 * SequenceAlignment.c:73).  acoss_dmax scores a zeroed table; the Serra09 call pattern
 * (qmax then dmax on the same table) is acoss_qmax_dmax.  Inputs too small for the
 * recurrence return 0.0 like the reference (SequenceAlignment.c:78-80,117-119,151-153).
 */
int acoss_qmax(acoss_ctx* ctx, const uint8_t* S, int32_t M, int32_t N, float* result);
int acoss_dmax(acoss_ctx* ctx, const uint8_t* S, int32_t M, int32_t N, float* result);
int acoss_qmax_dmax(acoss_ctx* ctx, const uint8_t* S, int32_t M, int32_t N, float* qmax_result, float* dmax_result);
int acoss_swconstrained(acoss_ctx* ctx, const uint8_t* S, int32_t N, int32_t M, float* result);

/*
 * The step before the path (Serra09.load_features, benchmarking/Serra09.py:96-114) on the
 * device: block aggregation of full-rate features over `fac` frames, slices
 * [fac*k, min(fac*(k+1), n)) as librosa.util.sync(data, arange(0, n, fac)) cuts them, with
 * aggregate 0 = np.median (chroma, Serra09.py:104) or 1 = np.mean (mfcc, :110), in the input's
 * precision and numpy's summation order; zero_nonfinite replaces NaN / Inf by 0 first
 * (Serra09.py:107-109).  Song s has nframes[s] frames at `full` + offsets[s]*bins elements,
 * layout 0 = frames x bins (hpcp / crema), 1 = bins x frames (mfcc_htk).  `out` receives
 * ceil(nframes/fac) x bins per song, frames x bins (what acoss_upload_songs takes), at
 * out + out_offsets[s]*bins elements.  gchroma_out (nullable, layout 0 only):
 * [n_songs][bins] global chroma sum_t / max (Serra09.py:24-28), same dtype.  Host pointers.
 */
int acoss_aggregate_songs(acoss_ctx* ctx, const void* full, int dtype, int layout,
                          const int64_t* offsets, const int32_t* nframes, int32_t n_songs,
                          int32_t bins, int32_t fac, int32_t aggregate, int32_t zero_nonfinite,
                          void* out, const int64_t* out_offsets, void* gchroma_out);

/*
 * The step after the path (CoverAlgorithm.getEvalStatistics, benchmarking/CoverAlgorithm.py:330-418)
 * on the device: for every song of a clique with at least two members, the 1-based positions of
 * the clique's OTHER songs in that song's row of the n x n score matrix D (float32, row-major, as
 * CoverAlgorithm.Ds[type]; the diagonal is ignored) sorted by descending score -- the reference's
 * `iranks[0:-1]` (:378-382).  Replaces np.argsort(-D, 1) (:362) and the O(n^2) Python loop
 * (:367-382); equal scores keep the clique-contiguous order (a stable sort; the reference's
 * introsort leaves it undefined).
 *   order        [n]     the clique-contiguous song order `idx` of :355 (cliques sorted by size,
 *                        descending), a permutation of 0..n-1
 *   rows                 number of leading positions that belong to cliques of >= 2 songs
 *   clique_start [rows]  position of the first song of the clique position p belongs to
 *   clique_size  [rows]  K of that clique
 *   out_offsets  [rows]  running sum of (K - 1): where row p's K - 1 ranks go in iranks_out
 *   iranks_out           ascending ranks, sum of (K - 1) int32
 * D is a host pointer, or a device pointer with ACOSS_FLAG_DEVICE_INPUT; the rest are host pointers.
 * MR / MRR / MDR / MAP / Top-K follow on the host from the integer ranks (float64, numpy's order).
 */
int acoss_eval_ranks(acoss_ctx* ctx, const float* D, int64_t n, const int32_t* order,
                     const int32_t* clique_start, const int32_t* clique_size, int32_t rows,
                     const int64_t* out_offsets, int32_t* iranks_out, uint32_t flags);
/* Device time (ms) of the k_eval_ranks launch of the last acoss_eval_ranks call. */
int acoss_last_eval_ms(acoss_ctx* ctx, float* out);

/*
 * Device-resident score matrices: after this call every acoss_serra09_pairs on the context ALSO writes, for each pair
 * (a, b) of its list, qmax_raw / (M + N) and dmax_raw / (M + N) -- what Serra09.similarity stores in
 * self.Ds[...][i][j] (benchmarking/Serra09.py:174-175; float64 division rounded to float32) -- to
 * qmax_dev[i * n + j] / dmax_dev[i * n + j], i = song_of_slot[a], j = song_of_slot[b] (identity when song_of_slot is
 * NULL; negative entries are skipped).  The matrices are the caller's device memory (n x n float32, e.g. torch tensors)
 * and stay the caller's; CoverAlgorithm.all_pairwise symmetrises them on the device and getEvalStatistics hands them to
 * acoss_eval_ranks with ACOSS_FLAG_DEVICE_INPUT, so the N x N matrix never crosses PCIe for the evaluation
 * (benchmarking/CoverAlgorithm.py:166-183, 345-382).  Either pointer may be NULL; both NULL unbinds.
 */
int acoss_bind_score_matrices(acoss_ctx* ctx, float* qmax_dev, float* dmax_dev, int64_t n, const int32_t* song_of_slot,
                              int32_t n_slots);

/* Test / tuning knobs (results never depend on them): "force_wide_sweep" (0/1),
 * "force_gscratch" (0/1), "dp_max_cpl" (4, 8, 16), "dp_packed" (0/1: Qmax + Dmax on packed 16-bit lanes, default 1),
 * "chunk_pairs" (pairs per internal chunk),
 * "sweep_gen" (2, 4 or 7: which generation of the cross-similarity sweep runs where it applies; default 7 = FP64
 * tensor-core sweep with chained row selection for column songs of up to 1015 frames, integer-key SIMT sweep (4) up to
 * 1320, float64 sweeps (2) beyond; 4 and 2 force the older kernels for cross-checks),
 * "eval_direct" (0/1, default 0: acoss_eval_ranks counts cliques of more than 5 songs with one compare per member
 * instead of the binary search + histogram kernel),
 * "tiled_wide" (0/1, default 1: column songs beyond 1408 frames go through the tiled float64 sweep,
 * kernels_sweep1t.cuh; 0 = the one-CTA general sweep they are checked against),
 * "gap_units" (2 .. 2^24: safety gap, in key units, below which a separator is re-decided in
 * float64; larger values only move work to the exact path), "single_stream" (0/1: keep the
 * long-song sweeps on the main stream), "debug_flags" (profiling experiments). */
int acoss_set_option(acoss_ctx* ctx, const char* name, int64_t value);

/* Profiling aid: counters of the gen-7 sweep, summed over all launches: [0] rows resolved, [1] rows decided in
 * float64, [2] strips, [3] items, [4..15] cycles per phase (kernels_sweep7.cuh).  All zero unless the library was
 * built with -DACOSS_S7_STATS. */
int acoss_debug_sweep7_stats(acoss_ctx* ctx, uint64_t out[16], int reset);

/* Measurement aid: out[0] = device time (ms, CUDA events) of the k_feat_aggregate launches of the last
 * acoss_aggregate_songs call, out[1] = the bytes they read and wrote (full-rate input + aggregated output). */
int acoss_last_aggregate_stats(acoss_ctx* ctx, double out[2]);

/* Profiling aid: host wall times (ms) of the last acoss_serra09_pairs call: [0] whole call, [1] preparing the first
 * chunk (not overlapped), [2] issuing copies + launches, [3] preparing chunks 2.. and handing out earlier results (overlapped with kernels),
 * [4] waiting for the stream, [5] handing results out (mostly inside [3]). */
int acoss_debug_host_ms(acoss_ctx* ctx, float out[6]);

/* Counters for bench.py: kernels launched by this context since creation / last reset. */
int64_t acoss_kernel_launches(acoss_ctx* ctx);
void acoss_reset_counters(acoss_ctx* ctx);

/* Device time (ms, CUDA events on the context's stream) spent in each stage during the
 * last acoss_serra09_pairs call: out[0]=oti, [1]=threshold sweeps, [2]=crp sweep, [3]=dp,
 * [4]=whole call on the stream.  Returns ACOSS_ERR_STATE before the first call. */
int acoss_last_stage_ms(acoss_ctx* ctx, float out[5]);

#ifdef __cplusplus
}
#endif
#endif /* ACOSS_B200_H */
