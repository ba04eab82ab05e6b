/*
 * melissa_b200 -- C ABI of the B200-native variational-Bayes hot path of Melissa.
 *
 * This is the drop-in boundary for the reference's `melissa()` /
 * `melissa_inner()` (reference R/melissa_vb.R:86-219 and :232-529).  The
 * reference is pure R and has no FFI of its own; these are the symbols an R
 * `.Call` / `.C` glue or any other host binds.  Plain
 * pointers and sizes only: no R, torch or C++ types cross this boundary, no
 * C++ exception leaves it.  Every function returns 0 on success and a
 * non-zero status on failure, with a message in melissa_last_error().
 *
 * Index conventions: everything is 0-based.  A "pair" is one covered
 * (cell, region) combination, i.e. one non-NA element X[[n]][[m]] of the
 * reference's nested input list (R/melissa_vb.R:11-17, :119-121).  Pairs are
 * passed cell-major (all pairs of cell 0, then cell 1, ...), regions ascending
 * inside a cell -- the order `which(!is.na(X[[n]]))` produces.
 * Array layouts on the way out are R's (column-major) so the glue can hand
 * the buffers to `array()` / `matrix()` unchanged.
 */
#ifndef MELISSA_B200_H
#define MELISSA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct melissa_handle melissa_handle;

/* status codes */
#define MELISSA_OK 0
#define MELISSA_ERR_ARG 1      /* bad argument (mirrors the reference's assert_that / stop) */
#define MELISSA_ERR_CUDA 2     /* CUDA runtime failure */
#define MELISSA_ERR_STATE 3    /* call out of order */

/* flags written by melissa_vb_run / melissa_vb_flags */
#define MELISSA_FLAG_LB_DECREASED 1   /* warning("Warning: Lower bound decreases!")  R/melissa_vb.R:495-497 */
#define MELISSA_FLAG_NOT_CONVERGED 2  /* warning("VB did not converge!")             R/melissa_vb.R:511-513 */

#define MELISSA_FLAG_INTERRUPTED 4    /* the progress callback asked to stop (see melissa_set_progress) */

/* basis kinds (BPRMeth basis objects; only the RBF object is pinned) */
#define MELISSA_BASIS_RBF 1

/* Message of the last failure on the calling thread ("" if none). */
const char *melissa_last_error(void);

/*
 * Progress / interrupt hook: the role of the reference's progress bar and its message every 50th iteration
 * (R/melissa_vb.R:305-307, :503-507), and the place where an R host checks for a user interrupt.  `fn` is called
 * on the CALLING thread at every lower bound of a run started from that thread (melissa_vb_run,
 * melissa_vb_run_multi) with the iteration, the bound and its change; a non-zero return stops the run after that
 * iteration: the results so far are returned and MELISSA_FLAG_INTERRUPTED is set.  The callback must not longjmp
 * (an R glue uses R_ToplevelExec around R_CheckUserInterrupt).  NULL removes the hook.
 */
typedef int (*melissa_progress_fn)(int32_t iteration, double lower_bound, double change, void *user);
int melissa_set_progress(melissa_progress_fn fn, void *user);

/* ABI version of the loaded library. */
int melissa_abi_version(void);

/*
 * Pack the observations once (replaces init_design_matrix / extract_y /
 * region_ind / cell_ind / HH construction: R/utils.R:64-97,
 * R/melissa_vb.R:119-131, :254-259).  Copies the host arrays to `device`,
 * builds the region-major index and the per-pair H^T H; the design matrices H
 * are never materialised (the RBF basis is re-evaluated in-kernel).
 *
 *   n_cells, n_regions   N and M of THIS shard (cells may be a shard of a larger job)
 *   n_pairs              P, number of covered (cell, region) pairs
 *   pair_ptr[P+1]        CpG offsets of each pair into x / y
 *   pair_cell[P]         cell of each pair, non-decreasing (cell-major)
 *   pair_region[P]       region of each pair, strictly increasing inside a cell
 *   x[nnz]               CpG locations scaled to [-1, 1] (column 1 of the region matrix)
 *   y[nnz]               methylation state, 0 or 1       (column 2)
 *   basis_kind           MELISSA_BASIS_RBF
 *   n_basis              basis$M (number of RBFs; D = n_basis + 1; 0 = mean-rate model)
 *   mus[n_basis], gamma  basis$mus, basis$gamma
 *
 * Errors mirror the reference: a cell without any covered region fails like
 * stop("No coverage across all regions. ...") (R/utils.R:69-71); a region no
 * cell covers is rejected (the reference's .add_func(NULL) would fail,
 * R/melissa_vb.R:368-370) unless allow_empty_regions != 0 (cell shards).
 */
int melissa_pack(int32_t n_cells, int32_t n_regions, int64_t n_pairs,
                 const int64_t *pair_ptr, const int32_t *pair_cell, const int32_t *pair_region,
                 const double *x, const uint8_t *y,
                 int32_t basis_kind, int32_t n_basis, const double *mus, double gamma,
                 int32_t device, int32_t allow_empty_regions, melissa_handle **out);

/*
 * Same as melissa_pack, but the five observation arrays are DEVICE pointers on
 * `device` (data generated or already resident there).  Nothing visits the host:
 * the index lists are validated by a kernel (same rules, same messages) and the
 * layout is built from them on the device.
 * No counterpart in the reference (its data always start as R lists).
 */
int melissa_pack_device(int32_t n_cells, int32_t n_regions, int64_t n_pairs,
                        const int64_t *pair_ptr, const int32_t *pair_cell, const int32_t *pair_region,
                        const double *x, const uint8_t *y,
                        int32_t basis_kind, int32_t n_basis, const double *mus, double gamma,
                        int32_t device, int32_t allow_empty_regions, melissa_handle **out);

/*
 * Library-wide tuning and diagnostic knobs; the defaults are what is measured and tested.  No counterpart in
 * the reference.
 * Read by the next melissa_pack*:
 *   "sliced"        -1 = the region-sliced layout whenever the RBF centres are equally spaced (default), 0 = always
 *                   the flat cell-major layout of round 1, 1 = require the sliced layout (pack fails otherwise)
 *   "cell_block"    cells per block of the sliced layout; 0 (default) = 256
 *   "pack_host"     1 = build the slice index with host threads instead of on the device (cross-check of the two builders)
 *   "sweep_chunk"   slices per chunk of the sweep's work list; 0 (default) = guided self-scheduling from the problem size
 *   "sweep_taper"   1 = the three-step chunk list (full / half / quarter size) instead of the guided one (comparison)
 *   "clusters_hint" accepted and ignored (round 1 chose the layout from it)
 * Read by the next launch:
 *   "sweep_group"   consecutive slices a warp claims at a time in long chunks (1..64; 0 = default 2); chunks shorter
 *                   than 640 slices are always claimed slice by slice
 *   "sweep_warps"   warps per CTA of the sweep; 0 (default) = as many of 20 as fit next to the block sums
 *   "contract_mma"  1 = FP64 tensor-core contraction (default; K <= 32 and D <= 8), 0 = CUDA-core contraction
 *   "contract_tma"  1 = its operand staged by cp.async.bulk + mbarrier (default), 0 = plain loads (bit-identical)
 *   "contract_warps" warps per region of the contraction (1..8; 0 = from the slices per region)
 *   "debug_estep"   1 = after every sweep a second pair of kernels recomputes every E-step sum from the stored records
 *                   and every H^T E[z] from x with library exp / erfcx, compares, and prints what differs to stderr
 *                   (the checker that stood in for compute-sanitizer; slow)
 */
int melissa_set_option(const char *key, int64_t value);

/* Release every device and host resource behind the handle (NULL is a no-op). */
int melissa_free(melissa_handle *h);

/*
 * One complete VB fit on one GPU: melissa_inner (R/melissa_vb.R:232-529).
 * All buffers are HOST memory owned by the caller.
 *
 *   K                  number of clusters
 *   w0[M*D*K]          initial means, R array dim c(M, D, K), column-major
 *   delta_0[K]         Dirichlet prior
 *   alpha_0, beta_0    Gamma prior; beta_0 = NaN means the reference default
 *                      sqrt(alpha_0 + M*D/2)                      (:271-273)
 *   vb_max_iter        the loop runs i = 2 .. vb_max_iter          (:309)
 *   epsilon_conv       stop when |LB - LB_prev| < epsilon_conv     (:499-501)
 *   lb_every           lower bound every lb_every-th i and at vb_max_iter (10 in the reference, :458)
 *  outputs (any may be NULL to skip):
 *   W[M*D*K]           posterior means, dim c(M, D, K)
 *   W_Sigma[K*M*D*D]   posterior covariances; block (k, m) at ((k*M)+m)*D*D, each D x D
 *   r_nk[N*K]          responsibilities, dim c(N, K) column-major
 *   delta, alpha, beta, pi_k   each [K]
 *   lb[lb_cap]         lower bounds; lb[0] = -Inf as in the reference (:242); NOT yet minus log(K!)
 *   n_lb               entries written to lb
 *   n_iter             last i executed
 *   flags              MELISSA_FLAG_* bits
 */
int melissa_vb_run(melissa_handle *h, int32_t K, const double *w0, const double *delta_0,
                   double alpha_0, double beta_0, int32_t vb_max_iter, double epsilon_conv,
                   int32_t lb_every,
                   double *W, double *W_Sigma, double *r_nk, double *delta, double *alpha,
                   double *beta, double *pi_k, double *lb, int32_t lb_cap, int32_t *n_lb,
                   int32_t *n_iter, int32_t *flags);

/*
 * Step-wise interface (what melissa_vb_run is built from).  A host that shards
 * cells over several GPUs (one process per GPU) drives these and all-reduces
 * the reduction buffer between the phases; see INTEGRATION.md.
 *
 *   melissa_vb_begin     state set-up + first E[z] sweep   (R/melissa_vb.R:262-302)
 *   melissa_vb_phase_a   responsibilities (:330-331) and this shard's
 *                        r-weighted sufficient statistics (:368-373 sums)
 *                        -> reduction buffer  [all-reduce SUM it here]
 *   melissa_vb_phase_b   q(w) solves, beta/delta/E[log pi] updates (:368-392,
 *                        :443-451), E[z] refresh sweep (:418-435); with_lb != 0
 *                        also accumulates this shard's lb_pz_qz (:469-473)
 *                        -> lb buffer            [all-reduce SUM it here]
 *   melissa_vb_lower_bound  assembles LB (:475-492), returns it to the host
 *
 * n_cells_total is N of the whole job (pi_k's denominator, :338).
 * `stream` is a cudaStream_t (NULL = default stream); the caller's collective
 * must be ordered on the same stream.
 */
int melissa_vb_begin(melissa_handle *h, int32_t K, const double *w0, const double *delta_0,
                     double alpha_0, double beta_0, int64_t n_cells_total, void *stream);
int melissa_vb_phase_a(melissa_handle *h);
int melissa_vb_phase_b(melissa_handle *h, int32_t with_lb);
int melissa_vb_lower_bound(melissa_handle *h, double *lb_out);

/* Device pointers + lengths (in doubles) of the two buffers a multi-GPU host all-reduces. */
int melissa_vb_reduce_buffer(melissa_handle *h, void **dev_ptr, int64_t *n_doubles);
int melissa_vb_lb_buffer(melissa_handle *h, void **dev_ptr, int64_t *n_doubles);

/*
 * Several GPUs of one box behind ONE call: what the reference's user gets from the single function call
 * melissa(...) (R/melissa_vb.R:86-90) and its no_cores knob (R/utils.R:30-43), for hosts (R's .Call, C,
 * Python) that run one process.  Cells are sharded over the GPUs in contiguous ranges balanced by CpG count;
 * the library runs one host thread and one stream per GPU; the only per-iteration exchange, the r-weighted
 * sufficient statistics (M K R + K + 1 doubles per GPU), is summed by a kernel on every GPU that reads the other
 * GPUs' partial sums over NVLink peer loads, in rank order (bit-identical on every GPU, no library collective).
 * These entry points visit every GPU from the calling thread and put its current device back before they return.
 *
 *   melissa_pack_multi    melissa_pack for the whole job: shards the HOST arrays, packs every shard on its GPU
 *                         (devices[g], or GPU g if devices is NULL), returns one object
 *   melissa_multi_create  the same object from handles the caller packed itself (one per shard, in cell order,
 *                         each with allow_empty_regions = 1); the handles stay the caller's
 *   melissa_vb_run_multi  melissa_vb_run on all shards; same arguments, outputs for the whole job
 *                         (r_nk has dim c(N_total, K))
 *   melissa_multi_free    releases the object (and the handles, if melissa_pack_multi made them)
 */
typedef struct melissa_multi melissa_multi;
int melissa_pack_multi(int32_t n_cells, int32_t n_regions, int64_t n_pairs,
                       const int64_t *pair_ptr, const int32_t *pair_cell, const int32_t *pair_region,
                       const double *x, const uint8_t *y,
                       int32_t basis_kind, int32_t n_basis, const double *mus, double gamma,
                       int32_t n_gpus, const int32_t *devices, melissa_multi **out);
int melissa_multi_create(int32_t n_gpus, melissa_handle **handles, melissa_multi **out);
int melissa_vb_run_multi(melissa_multi *mm, int32_t K, const double *w0, const double *delta_0,
                         double alpha_0, double beta_0, int32_t vb_max_iter, double epsilon_conv,
                         int32_t lb_every,
                         double *W, double *W_Sigma, double *r_nk, double *delta, double *alpha,
                         double *beta, double *pi_k, double *lb, int32_t lb_cap, int32_t *n_lb,
                         int32_t *n_iter, int32_t *flags);
int melissa_multi_free(melissa_multi *mm);
int32_t melissa_multi_gpus(const melissa_multi *mm);
/* Device time of the iteration loop of the last melissa_vb_run_multi (CUDA events on every GPU's stream from the
 * first iteration to the last, set-up and result fetch excluded; the maximum over the GPUs), and the kernels all
 * shards have launched so far: what bench.py reports for the single-process arm. */
double melissa_multi_last_loop_ms(const melissa_multi *mm);
int64_t melissa_multi_launch_count(const melissa_multi *mm);

/* Copy the current state to host buffers (same layouts as melissa_vb_run; NULL skips). */
int melissa_vb_fetch(melissa_handle *h, double *W, double *W_Sigma, double *r_nk,
                     double *delta, double *alpha, double *beta, double *pi_k);

/* Per-pair H^T E[z] of the last sweep, [P * D] row-major per pair (diagnostic / parity tests). */
int melissa_vb_fetch_HtEz(melissa_handle *h, double *out);

/*
 * Per-pair penalised probit MLE on the packed data: the role of
 * BPRMeth::infer_profiles_mle(model = "bernoulli", lambda = .5, opt_itnmax = 15) in the
 * reference's default initialisation (call site R/melissa_vb.R:136-145).  Newton
 * iterations on  sum_i log Phi(s_i h_i . w) - lambda |w|^2  for every covered pair.
 * w_out[P * D] is a HOST buffer, pair-major in the order the pairs were passed to
 * melissa_pack.  Needs the sliced layout.  (BPRMeth's optimiser is not in the reference
 * tree; parity with it is statistical)
 */
int melissa_profiles_mle(melissa_handle *h, double lambda, int32_t n_iter, double *w_out);
/* The same with w_out_device[P * D] in DEVICE memory on the handle's GPU, for a host that goes on clustering the
 * profiles there (the k-means of the default initialisation, R/melissa_vb.R:147-169). */
int melissa_profiles_mle_device(melissa_handle *h, double lambda, int32_t n_iter, double *w_out_device);

/* Block until all work queued on the handle's stream is done. */
int melissa_sync(melissa_handle *h);

/*
 * Optional per-kernel timing with CUDA events on the handle's stream (what
 * bench.py's roofline figure is measured with; the reference's only equivalent is
 * wrapping whole calls in microbenchmark, inst/examples/model-efficiency/).
 * melissa_profile_enable(h, 1) starts recording one event pair per launch,
 * (h, 0) stops and drops them; melissa_profile_read sums one kernel's launches.
 */
#define MELISSA_KERNEL_SWEEP 0     /* per-CpG E[z] sweep + next E-step terms */
#define MELISSA_KERNEL_RESP 1      /* responsibilities */
#define MELISSA_KERNEL_CONTRACT 2  /* r-weighted sufficient statistics */
#define MELISSA_KERNEL_SOLVE 3     /* batched D x D solves */
#define MELISSA_KERNEL_SCALARS 4   /* beta / delta / E[log pi] */
#define MELISSA_KERNEL_LB 5        /* lower-bound assembly */
int melissa_profile_enable(melissa_handle *h, int32_t on);
int melissa_profile_read(melissa_handle *h, int32_t kernel_id, double *total_ms, int64_t *count);

/* Number of kernels this handle has launched so far (bench.py's gpu_launches). */
int64_t melissa_launch_count(const melissa_handle *h);

/*
 * Shape of the packed layout (diagnostic; bench.py reports the row padding from it).  No
 * counterpart in the reference.  Writes up to 8 values: [0] 1 = region-sliced layout, 0 = flat;
 * [1] cells per block; [2] slices; [3] stored CpG slots (rows x 32 lanes, padding included);
 * [4] CpGs; [5] pairs; [6] chunks of the sweep; [7] warps per sweep CTA of the current fit.
 */
int melissa_layout_info(const melissa_handle *h, int64_t *out, int32_t n);

/*
 * Imputation pass (reference impute_test_met with basis = NULL, is_predictive =
 * TRUE: R/utils.R:225-284): pred[i] = round(sum_k r_nk[n,k] * Phi(h(x_i) . W[m,,k]), 3)
 * for held-out CpGs given as the same kind of cell-major pair list (cells
 * without held-out regions simply have no pair).  Stand-alone: takes the fitted
 * W (dim c(M, D, K)) and r_nk (dim c(N, K)) of a melissa object, column-major,
 * as HOST buffers; pred[nnz] is a HOST buffer.
 */
int melissa_impute(int32_t n_cells, int32_t n_regions, int32_t K, int32_t basis_kind, int32_t n_basis,
                   const double *mus, double gamma, const double *W, const double *r_nk,
                   int64_t n_pairs, const int64_t *pair_ptr, const int32_t *pair_cell,
                   const int32_t *pair_region, const double *x, double *pred, int32_t device);

/*
 * melissa_impute for hosts that cannot allocate device memory.  (No counterpart
 * in the reference: it exists because of how GPU processes are deployed, not
 * because of anything in Melissa.)
 *
 * Who needs it: a host process in which some other runtime has already claimed
 * all of the GPU's memory for its own pool, so that one more cudaMalloc fails; or
 * a service that imputes a small batch per request and wants neither the CUDA
 * allocator nor cudaFree's device-wide synchronisation in its request path.
 * melissa_impute allocates and frees seven device buffers on every call; this
 * variant calls no cudaMalloc / cudaFree and no allocating library routine.  Instead it stages its inputs and its output in
 * a fixed array of MELISSA_NOALLOC_SCRATCH_BYTES that is part of the library
 * image -- every GPU that loads the library holds that much memory until the
 * process exits, whether or not this function is ever called.
 *
 * Same arguments, same HOST buffers and the same result as melissa_impute.
 * Limits and rules:
 *   - the staged pieces (pair_ptr, pair_cell, pair_region, x, pred, W, r_nk, each
 *     rounded up to 256 bytes) must fit the array; a larger call is refused with
 *     MELISSA_ERR_ARG and a message that states both sizes -- split the batch or
 *     use melissa_impute;
 *   - the first call on a device looks up the array's device address
 *     (cudaGetSymbolAddress, once per device).  Make that first call before any
 *     stream capture starts.  The call copies and launches on the default stream
 *     and blocks the calling thread until the result is back;
 *   - pair_ptr[0] must be 0 (as for melissa_impute); the caller's current device is
 *     left as it was;
 *   - all calls share the one array and it keeps the previous call's bytes.  The
 *     function overwrites everything it reads before it launches, so stale
 *     contents are harmless, but two calls that are in flight at the same time on
 *     one GPU (two host threads) overwrite each other: serialise them.
 */
#define MELISSA_NOALLOC_SCRATCH_BYTES (8u << 20)   /* 8 MiB */
int melissa_impute_noalloc(int32_t n_cells, int32_t n_regions, int32_t K, int32_t basis_kind,
                           int32_t n_basis, const double *mus, double gamma, const double *W,
                           const double *r_nk, int64_t n_pairs, const int64_t *pair_ptr,
                           const int32_t *pair_cell, const int32_t *pair_region, const double *x,
                           double *pred, int32_t device);

#ifdef __cplusplus
}
#endif
#endif /* MELISSA_B200_H */
