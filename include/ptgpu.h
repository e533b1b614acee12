/*
 * include/ptgpu.h -- C ABI of the B200-native likelihood engine behind
 * pt::pll::Partition (matsengrp/pt, NNI-scoring hot path).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.
 * Every entry point names the libpll call it replaces in the reference's
 * wrapper (PTPLL = /root/reference/lib/libptpll/src, LIBPLL =
 * /root/reference/lib/libptpll/lib/pll-modules/libs/libpll/src).
 *
 * Index spaces are libpll's (LIBPLL/parse_utree.y:250-340) so host trees need
 * no change:  clv_index in [0, tips + clv_buffers)  (tips first),
 *             scaler_index in {-1} U [0, scale_buffers),
 *             matrix_index in [0, prob_matrices).
 * Extra clv/scale buffers (for directional CLVs, see ptgpu_score_nni_batch)
 * are simply requested at create time.
 *
 * Layouts in HBM (one engine = one alignment on one device):
 *   tips     uint8  [tips][sites]             4-bit state masks (A=1,C=2,G=4,T=8)
 *   clv      double [clv_buffers][sites][cats][4]
 *   scalers  uint32 [scale_buffers][sites]    counts of 2^256 rescalings
 *   pmatrix  double [prob_matrices][cats][4][4]   row = parent state
 *   sumtable double [sites][cats][4]
 *
 * Threading: a handle is used by one host thread at a time (as a
 * pll_partition_t is, PTPLL/pll_partition.hpp:26-33); many handles may be
 * used concurrently.  Each handle owns one CUDA stream.
 *
 * Errors: every function returns PTGPU_OK (0) or a negative code; a
 * thread-local message is available from ptgpu_last_error() (libpll:
 * PLL_FAILURE + thread-local pll_errno/pll_errmsg, LIBPLL/pll.c:24-25).
 * There is NO CPU fallback: without a CUDA device every call fails loudly.
 */
#ifndef PTGPU_H
#define PTGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTGPU_OK 0
#define PTGPU_ERR_CUDA (-1)      /* CUDA runtime failure (message has details) */
#define PTGPU_ERR_ARG (-2)       /* bad index / size / null pointer */
#define PTGPU_ERR_STATE (-3)     /* illegal tip state, model not set, ... */
#define PTGPU_ERR_NOT_FINITE (-4)/* d1/d2 not finite (PTPLL/pll_partition.cpp:443-449) */

#define PTGPU_SCALE_BUFFER_NONE (-1) /* == PLL_SCALE_BUFFER_NONE */

typedef struct ptgpu_engine ptgpu_engine;

/* Same field order and types as pll_operation_t (LIBPLL/pll.h:249-259), so a
   pll_operation_t array can be passed with a cast. */
typedef struct ptgpu_operation
{
  unsigned int parent_clv_index;
  int parent_scaler_index;
  unsigned int child1_clv_index;
  unsigned int child1_matrix_index;
  int child1_scaler_index;
  unsigned int child2_clv_index;
  unsigned int child2_matrix_index;
  int child2_scaler_index;
} ptgpu_operation_t;

/* One NNI candidate = the two CLV updates that face each other across the
   central edge after the swap, plus the start length of that edge.  The
   parent_* fields of p/q are ignored: the two new CLVs never leave the chip. */
typedef struct ptgpu_nni_candidate
{
  ptgpu_operation_t p;
  ptgpu_operation_t q;
  double length0;
} ptgpu_nni_candidate_t;

const char *ptgpu_last_error(void);
int ptgpu_device_count(void);

/* pll_partition_create (LIBPLL/pll.c:399-818) as called by
   PTPLL/pll_partition.cpp:47-57: 4 states, 1 rate matrix, per-site scalers. */
int ptgpu_create(unsigned int tips, unsigned int clv_buffers, unsigned int sites,
                 unsigned int rate_cats, unsigned int prob_matrices,
                 unsigned int scale_buffers, int device, ptgpu_engine **out);
void ptgpu_destroy(ptgpu_engine *e); /* pll_partition_destroy */

/* pll_set_frequencies + pll_set_subst_params + pll_set_category_rates
   (LIBPLL/models.c:366-400) followed by pll_update_eigen (:251-331).  The
   4x4 symmetric eigenproblem is solved on the host, once per model, and the
   result is kept as kernel constants.  subst = ac ag at cg ct gt. */
int ptgpu_set_model(ptgpu_engine *e, const double freqs[4], const double subst[6],
                    const double *category_rates);
/* pll_compute_gamma_cats(alpha, cats, rates, PLL_GAMMA_RATES_MEAN)
   (LIBPLL/gamma.c:220-292): host-only, needs no engine and no device. */
int ptgpu_gamma_rates_mean(double alpha, unsigned int cats, double *rates_out);
/* read back what the kernels use (test hook) */
int ptgpu_get_eigen(ptgpu_engine *e, double evecs[16], double inv_evecs[16],
                    double evals[4]);

/* pll_set_tip_states(partition, tip, pll_map_nt, sequence)
   (LIBPLL/pll.c:941-964, LIBPLL/maps.c:46-64): IUPAC nucleotide codes, gap/N/? =
   fully ambiguous; an unknown character is PTGPU_ERR_STATE.  `sequence` is a
   host buffer of `sites` characters; they are copied as they are and encoded
   to state masks by a kernel (pass pinned memory for an asynchronous copy). */
int ptgpu_set_tip_states(ptgpu_engine *e, unsigned int tip_index,
                         const char *sequence);

/* The loop of Partition::SetTipStates (lib/libptpll/src/pll_partition.cpp:160-199)
   in one call: sequences[i] (host, `sites` characters each) becomes tip
   tip_indices[i].  The raw characters go up back to back on the engine's
   stream (one copy when the sequences are contiguous in memory) and ONE kernel
   encodes them all; one synchronisation at the end instead of two per tip.
   An unknown character is PTGPU_ERR_STATE, as above. */
int ptgpu_set_tip_states_batch(ptgpu_engine *e, unsigned int count,
                               const unsigned int *tip_indices,
                               const char *const *sequences);

/* A second engine on the same alignment without a second upload: what a caller of the
   reference gets by constructing another pll_partition_t and repeating the per-tip
   pll_set_tip_states loop (lib/libptpll/src/pll_partition.cpp:160-199).  `dst` and `src`
   must have the same tip and site counts and live on the same device; the encoded tip
   states (and the pattern weights, if set) are copied device to device. */
int ptgpu_copy_tip_states(ptgpu_engine *dst, ptgpu_engine *src);

/* pll_set_pattern_weights (LIBPLL/pll.c:1047-1059): `weights` is a host array of
   `sites` multiplicities (every one >= 1: the reference's vector paths assume "no
   zero weights", LIBPLL/core_derivatives_avx.c:1123); NULL restores all-ones, the
   state after ptgpu_create and the only one pt itself uses.  From then on every sum
   over sites -- edge log-likelihoods, per-site log-likelihoods, both derivative sums,
   the scores of the batch entry points -- weights pattern i by weights[i], as
   libpll's kernels do.  CLVs, scalers and sumtables do not depend on the weights. */
int ptgpu_set_pattern_weights(ptgpu_engine *e, const unsigned int *weights);

/* pll_compress_site_patterns (LIBPLL/compress.c:138-282) for nucleotide data (pll_map_nt):
   the unique columns of `count` sequences of *length characters, in the reference's
   order (ascending by the columns' state codes, first taxon most significant), written
   over the front of each sequence (NUL-terminated there), *length set to their number
   and their multiplicities stored in weights_out[0 .. *length).  Characters come back
   as the canonical symbol of their state code, as in the reference.  Host-only: the
   alignment is compressed before it is uploaded.  An unknown character is
   PTGPU_ERR_STATE and leaves the sequences untouched. */
int ptgpu_compress_site_patterns(char **sequences, unsigned int count, unsigned int *length,
                                 unsigned int *weights_out);

/* pll_update_prob_matrices (LIBPLL/models.c:333-364) */
int ptgpu_update_prob_matrices(ptgpu_engine *e, const unsigned int *matrix_indices,
                               const double *branch_lengths, unsigned int count);

/* pll_update_partials (LIBPLL/partials.c:177-213): ops run in list order. */
int ptgpu_update_partials(ptgpu_engine *e, const ptgpu_operation_t *ops,
                          unsigned int count);

/* pll_compute_edge_loglikelihood (LIBPLL/likelihood.c:478-513); persite_lnl
   may be NULL, else `sites` doubles on the host. */
int ptgpu_compute_edge_loglikelihood(ptgpu_engine *e, unsigned int parent_clv_index,
                                     int parent_scaler_index,
                                     unsigned int child_clv_index,
                                     int child_scaler_index,
                                     unsigned int matrix_index,
                                     double *persite_lnl, double *lnl_out);

/* NO COUNTERPART IN THE REFERENCE: added for callers
   that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
   that forbid it).  The same computation as ptgpu_compute_edge_loglikelihood (without the
   per-site output) by the same kernel, but its scratch -- one partial per tile of 256 sites,
   the arrival counter, the result -- lives in a file-scope `static __device__` array of the
   library instead of engine-allocated buffers, and the call makes no cudaMalloc / cudaFree /
   allocating library call.  Limit: engines of at most ptgpu_noalloc_max_sites() = 2,097,152
   sites (8192 tiles; the array is 64 KiB and every GPU that loads the library holds it until
   exit); a larger engine gets PTGPU_ERR_ARG, a site shard PTGPU_ERR_STATE.  ALL engines on a
   device share the one array and it keeps the last call's contents (the call clears the counter
   word, everything else is overwritten before it is read): calls that overlap on the same GPU
   race -- issue them one after the other.  The result equals the other entry point's bit for
   bit (same reduction tree). */
unsigned int ptgpu_noalloc_max_sites(void);
int ptgpu_compute_edge_loglikelihood_noalloc(ptgpu_engine *e, unsigned int parent_clv_index,
                                             int parent_scaler_index,
                                             unsigned int child_clv_index,
                                             int child_scaler_index, unsigned int matrix_index,
                                             double *lnl_out);

/* pll_update_sumtable (LIBPLL/derivatives.c:164-234) into the engine-owned
   device sumtable (PTPLL keeps one scratch sumtable per partition too). */
int ptgpu_update_sumtable(ptgpu_engine *e, unsigned int parent_clv_index,
                          unsigned int child_clv_index, int parent_scaler_index,
                          int child_scaler_index);

/* pll_compute_likelihood_derivatives (LIBPLL/derivatives.c:243-312): first
   and second derivative of -lnL at branch_length from the current sumtable. */
int ptgpu_compute_likelihood_derivatives(ptgpu_engine *e, double branch_length,
                                         double *d_f, double *dd_f);

/* The body of Partition::OptimizeBranch (PTPLL/pll_partition.cpp:408-488)
   as one device-resident loop: sumtable, <= max_iter Newton steps with the
   reference's stopping rules, then the edge's P matrix at the new length.
   map_lambda = 0 outside MAP mode (10 in MAP mode).  *iterations (optional)
   receives the number of derivative evaluations. */
int ptgpu_optimize_branch(ptgpu_engine *e, unsigned int parent_clv_index,
                          int parent_scaler_index, unsigned int child_clv_index,
                          int child_scaler_index, unsigned int matrix_index,
                          double length0, unsigned int max_iter, double epsilon,
                          double map_lambda, double *length_out, int *iterations);

/* One edge of pllmod's local branch-length optimisation, i.e. the body of
   recomp_iterative up to and including pllmod_opt_minimize_newton
   (PLLMOD/optimize/pll_optimize.c:771-827, opt_algorithms.c:71-172): sumtable
   of the edge, then the safeguarded Newton-Raphson / bisection iteration in
   [xmin, xmax] from xguess with the reference's stopping rules, all in one
   device-resident loop.  PLLMOD = /root/reference/lib/libptpll/lib/pll-modules/src.
   *status: 0 converged, PTGPU_NEWTON_DERIV (derivative not finite; the
   reference sets PLLMOD_OPT_ERROR_NEWTON_DERIV and returns -inf),
   PTGPU_NEWTON_LIMIT (max_iter evaluations used up; the reference sets
   PLLMOD_OPT_ERROR_NEWTON_LIMIT).  No P matrix is touched: the caller decides
   (keep_update, pll_optimize.c:829-868). */
#define PTGPU_NEWTON_DERIV 1
#define PTGPU_NEWTON_LIMIT 2
int ptgpu_optimize_branch_bounded(ptgpu_engine *e, unsigned int parent_clv_index,
                                  int parent_scaler_index, unsigned int child_clv_index,
                                  int child_scaler_index, double xguess, double xmin,
                                  double xmax, double tolerance, unsigned int max_iter,
                                  double *length_out, int *status, int *iterations);

/* Batched SingleBranchOptimizer::EvaluateMove
   (src/move_tester/single_branch_optimizer.cpp:23-64) for `count` candidates
   of one parent tree: for each, the two CLV updates across the swapped edge,
   the sumtable, Newton on the central branch from length0, and the
   log-likelihood at the optimum.  Nothing in the engine's CLV / scaler /
   P-matrix state is modified.  status_out[i] is PTGPU_OK or
   PTGPU_ERR_NOT_FINITE. */
int ptgpu_score_nni_batch(ptgpu_engine *e, const ptgpu_nni_candidate_t *cands,
                          unsigned int count, unsigned int max_iter,
                          double epsilon, double map_lambda, double *length_out,
                          double *lnl_out, int *status_out);

/* Batched BranchNeighborhoodOptimizer::EvaluateMove
   (src/move_tester/branch_neighborhood_optimizer.cpp:27-80) for `count`
   candidates of one parent tree: per candidate the CLV updates across the
   swapped edge, then pllmod_opt_optimize_branch_lengths_local
   (PLLMOD/optimize/pll_optimize.c:969-1091) and the final edge
   log-likelihood, all candidates in one persistent kernel.  The caller
   flattens recomp_iterative's recursion (pll_optimize.c:771-934) into a step
   list of the same shape for every candidate: `n_prologue` steps run once,
   then `n_body` steps per smoothing, given TWICE: rows [n_prologue,
   n_prologue + n_body) are the first smoothing (nodes the recursion has not
   re-oriented yet are still read from the engine's CLVs), rows
   [n_prologue + n_body, n_prologue + 2 n_body) every later one (all touched
   nodes are private copies by then).  steps[k * count + i] is step k of
   candidate i and step_types[k] (n_prologue + 2 n_body entries) its type for
   all candidates (a candidate without that step has PTGPU_BLO_NOP there).
     PTGPU_BLO_OP : CLV a <- (CLV b through matrix b) x (CLV c through matrix c)
     PTGPU_BLO_OPT: optimise the edge between CLV a ("parent" of the sumtable,
                    tr_p) and CLV b (tr_p->back) whose P matrix is b_matrix and
                    whose length lives in length_slot: safeguarded Newton in
                    [bl_min, bl_max], tolerance bl_min/10, <= 30 evaluations
   Nothing in the engine's own CLV / scaler / P-matrix state is modified:
   candidate i writes only PRIVATE slots, addressed above the engine's ranges:
     CLV     tips + clv_buffers + i * n_private_clvs + j   (scaler: scale_buffers + same offset)
     matrix  prob_matrices + i * n_lengths + slot          (initialised from lengths0[i][slot])
   smooth_edges[i] is the edge whose lnL ends every smoothing (parent tree->back,
   child tree; pll_optimize.c:1049-1056), final_edges[i] the edge of the final
   Partition::LogLikelihood(node).  status_out[i]: 0, or PTGPU_NEWTON_DERIV /
   PTGPU_NEWTON_LIMIT if a solve failed (the reference stops optimising that
   candidate; its lnL is then taken on the state it was left in). */
#define PTGPU_BLO_NOP 0
#define PTGPU_BLO_OP 1
#define PTGPU_BLO_OPT 2
#define PTGPU_BLO_MAT 3 /* ptgpu_optimize_all_branches only: P matrix b_matrix at the length in length_slot */
typedef struct ptgpu_blo_step
{
  int type;
  unsigned int a_clv;
  int a_scaler;
  unsigned int b_clv, b_matrix;
  int b_scaler;
  unsigned int c_clv, c_matrix;
  int c_scaler;
  unsigned int length_slot;
} ptgpu_blo_step_t;
typedef struct ptgpu_blo_edge
{
  unsigned int parent_clv;
  int parent_scaler;
  unsigned int child_clv;
  int child_scaler;
  unsigned int matrix;
  unsigned int reserved;
} ptgpu_blo_edge_t;
int ptgpu_optimize_local_batch(ptgpu_engine *e, unsigned int count, unsigned int n_private_clvs,
                               unsigned int n_lengths, unsigned int n_prologue, unsigned int n_body,
                               const int *step_types, const ptgpu_blo_step_t *steps,
                               const ptgpu_blo_edge_t *smooth_edges,
                               const ptgpu_blo_edge_t *final_edges, const double *lengths0,
                               double bl_min, double bl_max, double tolerance,
                               unsigned int smoothings, double *lengths_out, double *lnl_out,
                               int *status_out, int *smoothings_out);

/* Partition::OptimizeAllBranches (lib/libptpll/src/pll_partition.cpp:542-582, the
   non-pllmod branch) as ONE persistent kernel instead of ~10 engine calls per
   branch and pass.  The reference's procedure is: orient the CLVs towards the
   tree's node, lnL; then passes of OptimizeAllBranchesOnce (:490-540: for every
   node in post-order a PARTIAL TraversalUpdate towards it and OptimizeBranch,
   :408-488) each followed by a PARTIAL traversal back and lnL, while the lnL
   still moves by more than `epsilon` (at most 1 + max_passes - 1 passes).
   Which CLVs and matrices each traversal recomputes depends only on the
   orientation flags, so the caller (who owns the flags) unrolls the procedure
   into steps on the ENGINE'S OWN buffers -- unlike ptgpu_optimize_local_batch
   nothing is private here, the engine ends in the state the call sequence
   would have left:
     PTGPU_BLO_MAT  matrix b_matrix := P(length of edge length_slot); b_matrix == length_slot
     PTGPU_BLO_OP   CLV update, as ptgpu_update_partials
     PTGPU_BLO_OPT  OptimizeBranch on the edge (a = node, b = node->back, matrix
                    b_matrix == length_slot): sumtable, <= max_iter Newton steps
                    with the reference's stopping rules and exponential prior
                    rate map_lambda (0 = ML), new length into the slot, matrix
                    recomputed
   steps[0 .. n_prologue) run once (the first traversal); then pass k runs rows
   n_prologue + min(k, n_bodies-1) * n_body .. + n_body (the traversals of a pass
   depend on the orientation the previous pass left behind; from some pass on
   they repeat).  Length slots are the engine's matrix indices; lengths0 /
   lengths_out have prob_matrices entries.  lnl_edge is the edge of
   Partition::LogLikelihood(tree); with n_prior > 0 every lnL also gets the MAP
   prior terms of the listed edges, added in that order.  *lnl_out is the last
   lnL, *passes_out the passes run; *status_out 0, 1 (d1 not finite) or 2 (d2),
   in which case PTGPU_ERR_NOT_FINITE is returned as well. */
int ptgpu_optimize_all_branches(ptgpu_engine *e, unsigned int n_prologue, unsigned int n_body,
                                unsigned int n_bodies, const ptgpu_blo_step_t *steps,
                                const ptgpu_blo_edge_t *lnl_edge, const double *lengths0,
                                const unsigned int *prior_edges, unsigned int n_prior,
                                unsigned int max_iter, double epsilon, double map_lambda,
                                unsigned int max_passes, double *lengths_out, double *lnl_out,
                                int *status_out, int *passes_out);

/* ---- site shards: one engine per GPU holds a contiguous range
   of the alignment's sites.  CLV updates, P matrices and sumtables are purely
   local; every reduction over sites (edge lnL; d1/d2 of each Newton step; the
   per-candidate lnL of a batch) is formed locally in a fixed order and then
   summed over shards by the caller's hook before the scalar update -- the
   scheme of pll-modules' parallel_reduce_cb
   (PLLMOD/optimize/pll_optimize.c:1211-1212 lnL, :1249-1255 df/ddf).
   The hook must sum `count` doubles at `device_buf` IN PLACE over all shards,
   ordered on `cuda_stream` (e.g. ncclAllReduce(ncclDouble, ncclSum) on that
   stream), and return 0.  All shards make the same sequence of calls. */
typedef int (*ptgpu_allreduce_fn)(void *ctx, double *device_buf, size_t count,
                                  void *cuda_stream);
int ptgpu_set_allreduce(ptgpu_engine *e, ptgpu_allreduce_fn fn, void *ctx);
/* where this engine's sites sit in the whole alignment (keeps the reference's
   4-site vector / scalar tail split of the derivative loop at the GLOBAL
   position, LIBPLL/core_derivatives_avx.c:1111-1156) */
int ptgpu_set_shard(ptgpu_engine *e, unsigned long long site_offset,
                    unsigned long long global_sites);

/* ---- site shards, the engines' own exchange (SURVEY.md 8b "sharded: ptgpu_comm_init",
   8e): instead of a caller-supplied hook the kernels sum the per-tree scalars themselves,
   over NVLink peer memory -- every shard owns a small mailbox that all shards have mapped;
   a value is published with plain stores into every shard's mailbox and collected by
   polling the local one; sums are formed in rank order, so every shard sees the same bits
   and results are reproducible per shard count.  With this exchange in place the
   step-program entry points (ptgpu_optimize_all_branches, ptgpu_optimize_local_batch) run
   on shards as ONE persistent kernel per GPU: no launch, no host round trip and no
   collective call between the Newton steps of a whole-tree optimisation.  It replaces what
   a multi-partition caller of pll-modules would supply as parallel_reduce_cb
   (PLLMOD/optimize/pll_optimize.c:1211-1212 lnL, :1249-1255 df/ddf).
   All shards must make the same sequence of engine calls (they compute the same sums, so
   they take the same decisions); a shard whose peer does not answer within
   PTGPU_SHARD_TIMEOUT_MS (default 20000) fails the call with PTGPU_ERR_STATE instead of
   hanging.  At most 8 shards.

   One process drives all shards (what pt::pll::Partition does with PTGPU_SITE_SHARDS=n):
   shards[r] becomes rank r; the engines may live on different devices (peer access is
   enabled) or, for tests, on the same one.  Call ptgpu_set_shard on each as well. */
int ptgpu_comm_init(ptgpu_engine *const *shards, unsigned int nshards);
/* One process per shard (torchrun): every rank exports a handle to its mailbox
   (a cudaIpcMemHandle_t), the ranks exchange the handles by whatever means they have
   (e.g. torch.distributed.all_gather_object), then every rank connects with the full
   list (handles[r] = rank r's; its own entry is ignored). */
typedef struct ptgpu_shard_handle { unsigned char bytes[64]; } ptgpu_shard_handle_t;
int ptgpu_shard_export(ptgpu_engine *e, ptgpu_shard_handle_t *out);
int ptgpu_shard_connect(ptgpu_engine *e, unsigned int rank, unsigned int nranks,
                        const ptgpu_shard_handle_t *handles);
int ptgpu_shard_disconnect(ptgpu_engine *e);
/* shard count (0 = not connected), the sticky time-out flag and the number of exchanges
   completed since connect (synchronises the engine's stream) */
int ptgpu_shard_status(ptgpu_engine *e, int *nranks_out, int *error_out,
                       unsigned long long *exchanges_out);

/* test / debug hooks: host copies of device state */
int ptgpu_get_clv(ptgpu_engine *e, unsigned int clv_index, double *out);
int ptgpu_set_clv(ptgpu_engine *e, unsigned int clv_index, const double *in);
int ptgpu_get_scaler(ptgpu_engine *e, int scaler_index, unsigned int *out);
int ptgpu_get_pmatrix(ptgpu_engine *e, unsigned int matrix_index, double *out);
int ptgpu_set_pmatrix(ptgpu_engine *e, unsigned int matrix_index, const double *in);
int ptgpu_get_sumtable(ptgpu_engine *e, double *out);

/* doubles per site of the sumtables ptgpu_score_nni_batch keeps per candidate: 4 x rate_cats
   (the layout of pll_update_sumtable), or 3 x rate_cats + 1 when the model has a zero
   eigenvalue (every reversible model) -- that column multiplies exp(0) = 1 in every Newton
   evaluation (LIBPLL/core_derivatives.c:560-575) and is folded into one double per site.
   What bench.py's algorithmic byte count of a Newton sweep is based on. */
unsigned int ptgpu_batch_sumtable_doubles(ptgpu_engine *e);

/* counters since create: [0] kernel launches, [1] CLV site-updates,
   [2] Newton derivative passes (site sweeps), [3] NNI candidates scored,
   [4] batch sweeps that also summed the site log-likelihoods (they read the
   scaler word too), [5] extra lnL-only sweeps after the Newton loop */
int ptgpu_get_counters(ptgpu_engine *e, unsigned long long out[6]);

/* Optional per-kernel timing with CUDA events on the engine's stream (used by
   bench.py for the roofline figures; off by default).  Kinds index the arrays
   of ptgpu_get_kernel_times, which drains pending events (synchronises). */
#define PTGPU_KERNEL_PMATRIX 0
#define PTGPU_KERNEL_PARTIALS 1
#define PTGPU_KERNEL_EDGE_LNL 2
#define PTGPU_KERNEL_SUMTABLE 3
#define PTGPU_KERNEL_NEWTON 4
#define PTGPU_KERNEL_NNI_PREPARE 5
#define PTGPU_KERNEL_TIP_ENCODE 6
#define PTGPU_KERNEL_KINDS 7
int ptgpu_set_profiling(ptgpu_engine *e, int on);
int ptgpu_get_kernel_times(ptgpu_engine *e, double ms_out[PTGPU_KERNEL_KINDS],
                           unsigned long long launches_out[PTGPU_KERNEL_KINDS], int reset);

/* the engine's stream (a cudaStream_t) so callers can order their own CUDA
   events / work against it */
void *ptgpu_stream(ptgpu_engine *e);
int ptgpu_synchronize(ptgpu_engine *e);

#ifdef __cplusplus
}
#endif
#endif /* PTGPU_H */
