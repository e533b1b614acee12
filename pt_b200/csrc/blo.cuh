// pt_b200/csrc/blo.cuh -- batched local branch-length optimisation on the device.
//
// BranchNeighborhoodOptimizer::EvaluateMove
// (/root/reference/src/move_tester/branch_neighborhood_optimizer.cpp:27-80) scores
// ONE NNI candidate with ~100 libpll calls: two CLV updates, then pll-modules'
// pllmod_opt_optimize_branch_lengths_local (PLLMOD/optimize/pll_optimize.c:969-1091:
// up to 32 smoothings of recomp_iterative, :771-934 -- sumtable, safeguarded
// Newton, P matrix, re-orient the node's CLV, recurse), an edge lnL per smoothing,
// and finally a FULL traversal to undo the damage.  Here ALL candidates of a tree
// run that procedure at once in one persistent cooperative kernel:
//
//  * the host flattens recomp_iterative's recursion for the requested radius into
//    a short step list per candidate (CLV update / optimise edge), identical in
//    shape for all candidates (missing subtrees become no-ops);
//  * every candidate owns private CLV, scaler, P-matrix and length slots, so the
//    engine's own state is never touched (no resynchronising traversal);
//  * the kernel walks the step list in lock step for all live candidates, each
//    phase spread over the whole grid as (candidate, tile) work items, with the
//    same elementwise arithmetic and the same fixed-order reductions as the
//    single-edge kernels -- a candidate's numbers do not depend on its batch.
#pragma once

#include "kernels.cuh"

namespace ptgpu {

// index spaces seen by the batch: [0, *_total) are the engine's, above that the
// candidates' private slots
struct BloView
{
  DevState s;
  unsigned int clv_total;   // tips + clv_buffers
  unsigned int scal_total;  // scale_buffers
  unsigned int mat_total;   // prob_matrices
  double *pclv;             // [private clvs][clv_stride]
  unsigned int *pscal;      // [private scalers][scaler_stride]
  double *pmat;             // [private matrices][C][16]
};

enum : int { BLO_NOP = 0, BLO_OP = 1, BLO_OPT = 2, BLO_MAT = 3 };

struct BloStep
{
  int type;
  // BLO_OP : parent (a) <- child1 (b via b_mat) x child2 (c via c_mat)
  // BLO_OPT: edge with "parent" side a and "child" side b (sumtable order of
  //          recomp_iterative: tr_p, tr_p->back), its P matrix b_mat and length slot
  unsigned int a_clv;
  int a_scal;
  unsigned int b_clv, b_mat;
  int b_scal;
  unsigned int c_clv, c_mat;
  int c_scal;
  unsigned int slot;
};

struct BloEdge  // an edge whose log-likelihood is taken: parent side, child side, matrix
{
  unsigned int p_clv;
  int p_scal;
  unsigned int c_clv;
  int c_scal;
  unsigned int mat;
  unsigned int pad;
};

struct BloCand
{
  double lnl;        // lnL after the last completed smoothing (start: before the first)
  double final_lnl;  // LogLikelihood(node) at the end
  int active;        // still smoothing
  int failed;        // a Newton solve failed (pll_errno set in the reference): no further steps
  int smoothings;    // completed smoothings
  int newton_evals;  // statistics
};

struct BloArgs
{
  BloView v;
  ModelConst m;
  const BloStep *steps;      // [n_pro + 2 n_body][B]: prologue, first smoothing, later smoothings
  const BloEdge *smooth_edge; // [B] parent q, child p  (pll_optimize.c:1000-1007,1049-1056)
  const BloEdge *final_edge;  // [B] parent p, child q  (Partition::LogLikelihood(node))
  BloCand *cand;             // [B]
  double *lens;              // [B][NL] private branch lengths
  NewtonState *nst;          // [2][B]: Newton state, double-buffered by evaluation parity
  int *fin;                  // [B]: the current OPT step's solve of this candidate has finished
  double *sumtables;         // [B][entry_stride]
  size_t entry_stride;
  double *partials;          // [2][B][nchunks][3] Newton chunk partials, double-buffered likewise
  double *tile_partials;     // [B][ntiles] edge-lnL tile partials
  const int *step_type;      // [n_pro + 2 n_body] the (uniform) type of each step position
  unsigned int B, NL, n_pro, n_body;
  unsigned int n_bodies;     // body variants: smoothing k runs variant min(k, n_bodies - 1)
  unsigned int sites, ntiles, chunk_tiles, nchunks;
  unsigned int smoothings, max_newton;
  double xmin, xmax, xtol, tolerance;
  unsigned long long vec_end;
  // what the procedure is:
  //  0  pllmod_opt_optimize_branch_lengths_local (PLLMOD/optimize/pll_optimize.c:969-1091):
  //     safeguarded Newton per edge (:771-827), stop on a worse or converged lnL
  //  1  pt's Partition::OptimizeAllBranches (PTPLL/pll_partition.cpp:542-582): plain Newton of
  //     Partition::OptimizeBranch (:408-488), stop when |lnL change| <= tolerance
  int procedure;
  double lambda;             // MAP prior rate (procedure 1), 0 = ML
  const unsigned int *prior_slots;  // length slots whose prior terms LogLikelihood adds, in its order
  unsigned int n_prior;
  // site shards (kernels.cuh, ShardComm): this engine holds one contiguous range of the
  // alignment's sites; every sum over sites (d1/d2 of a Newton evaluation, an edge lnL) is
  // formed locally in the usual fixed order and then summed over the shards in rank order,
  // inside the kernel, over peer memory.  All shards run the same program on the same sums,
  // so they take the same decisions and stay in lock step.  nranks <= 1: not sharded.
  ShardComm comm;
};

// ---- loaders over the combined index spaces ---------------------------------
// Private CLVs, scalers and matrices are rewritten between phases by other SMs:
// all their reads go to L2 (ld.global.cg), never through L1 or the read-only path.
__device__ __forceinline__ void ld256_cg(const double *p, double v[4])
{
  asm volatile("ld.global.cg.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
               : "l"(p));
}

template <int C>
__device__ __forceinline__ const double *blo_clv_ptr(const BloView &v, unsigned int idx)
{
  return idx < v.clv_total ? v.s.clv + (size_t)(idx - v.s.n_tips) * v.s.clv_stride
                           : v.pclv + (size_t)(idx - v.clv_total) * v.s.clv_stride;
}

template <int C>
__device__ __forceinline__ void blo_load_entry(const BloView &v, unsigned int idx, unsigned int site,
                                               int cat, double x[4])
{
  if (idx < v.s.n_tips)
  {
    const unsigned int mk = v.s.tips[(size_t)idx * v.s.tip_stride + site];
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = ((mk >> j) & 1u) ? 1.0 : 0.0;
  }
  else
    ld256_cg(blo_clv_ptr<C>(v, idx) + ((size_t)site * C + cat) * 4, x);
}

__device__ __forceinline__ unsigned int *blo_scal_ptr(const BloView &v, int idx)
{
  return (unsigned int)idx < v.scal_total ? v.s.scalers + (size_t)idx * v.s.scaler_stride
                                          : v.pscal + (size_t)(idx - v.scal_total) * v.s.scaler_stride;
}

__device__ __forceinline__ unsigned int blo_load_scaler(const BloView &v, int idx, unsigned int site)
{
  return idx < 0 ? 0u : __ldcg(blo_scal_ptr(v, idx) + site);
}

template <int C>
__device__ __forceinline__ double *blo_mat_ptr(const BloView &v, unsigned int idx)
{
  return idx < v.mat_total ? v.s.pmat + (size_t)idx * C * 16
                           : v.pmat + (size_t)(idx - v.mat_total) * C * 16;
}

template <int C>
__device__ __forceinline__ void blo_load_matrix(const BloView &v, unsigned int idx, int cat, double m[16])
{
  const double *pm = blo_mat_ptr<C>(v, idx) + cat * 16;
#pragma unroll
  for (int i = 0; i < 4; ++i) ld256_cg(pm + 4 * i, m + 4 * i);
}

// ---- tile bodies (T = kSitesPerPass*C threads, tile = kTileSites sites) --------

// The P matrices a work item needs (all C categories of up to two matrices), staged once
// per work item in shared memory and consumed row by row (as partials_kernel_v4): a thread
// then holds one 4-double row at a time instead of two 16-double matrices.
template <int C>
__device__ __forceinline__ void blo_stage_matrices(const BloView &v, unsigned int m0, unsigned int m1,
                                                   int count, double *smat)
{
  __syncthreads();   // the previous work item's rows are no longer read
  for (int i = threadIdx.x; i < count * C * 16; i += kSitesPerPass * C)
  {
    const int which = i / (C * 16), c = (i / 16) % C, k = i % 16;
    smat[(which * C + c) * kMatPad + k] = __ldcg(blo_mat_ptr<C>(v, which ? m1 : m0) + c * 16 + k);
  }
  __syncthreads();
}

// one CLV update on one tile: LIBPLL/core_partials_avx.c:366-529 (as partials_kernel_v4)
template <int C>
__device__ __forceinline__ void blo_op_tile(const BloView &v, const BloStep &st, unsigned int tile,
                                            const double *smat)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  const double *lmat = smat + cat * kMatPad;
  const double *rmat = smat + (C + cat) * kMatPad;
  double *dst = const_cast<double *>(blo_clv_ptr<C>(v, st.a_clv));
  unsigned int *sdst = st.a_scal >= 0 ? blo_scal_ptr(v, st.a_scal) : nullptr;
#pragma unroll 1
  for (int half = 0; half < kPasses; half += 2)
  {
    double l[2][4], r[2][4], p[2][4];
    unsigned int sc[2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = min(site0 + (half + j) * kSitesPerPass, v.s.sites - 1);
      blo_load_entry<C>(v, st.b_clv, site, cat, l[j]);
      blo_load_entry<C>(v, st.c_clv, site, cat, r[j]);
      sc[j] = blo_load_scaler(v, st.b_scal, site) + blo_load_scaler(v, st.c_scal, site);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
      double lrow[4], rrow[4];
      lds256(lmat + 4 * i, lrow);
      lds256(rmat + 4 * i, rrow);
#pragma unroll
      for (int j = 0; j < 2; ++j) p[j][i] = __dmul_rn(dot4(lrow, l[j]), dot4(rrow, r[j]));
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = site0 + (half + j) * kSitesPerPass;
      bool small = (p[j][0] < PT_SCALE_THRESHOLD) & (p[j][1] < PT_SCALE_THRESHOLD) &
                   (p[j][2] < PT_SCALE_THRESHOLD) & (p[j][3] < PT_SCALE_THRESHOLD);
      small = site_all<C>(small) && sdst != nullptr;
      if (small)
      {
#pragma unroll
        for (int i = 0; i < 4; ++i) p[j][i] = __dmul_rn(p[j][i], PT_SCALE_FACTOR);
      }
      if (site < v.s.sites)
      {
        st256(dst + ((size_t)site * C + cat) * 4, p[j]);
        if (cat == 0 && sdst) sdst[site] = sc[j] + (small ? 1u : 0u);
      }
    }
  }
}

// sumtable of one edge on one tile (as sumtable_kernel)
template <int C>
__device__ __forceinline__ void blo_sumtable_tile(const BloView &v, const ModelConst &m,
                                                  unsigned int pclv, unsigned int cclv,
                                                  double *sumtable, unsigned int tile)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  double p[kPasses][4], c[kPasses][4];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = min(site0 + ps * kSitesPerPass, v.s.sites - 1);
    blo_load_entry<C>(v, pclv, site, cat, p[ps]);
    blo_load_entry<C>(v, cclv, site, cat, c[ps]);
  }
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = site0 + ps * kSitesPerPass;
    double out[4];
    sumtable_entry<C>(m, p[ps], c[ps], out);
    if (site < v.s.sites) st256(sumtable + ((size_t)site * C + cat) * 4, out);
  }
}

// sumtable of one tile AND its contribution to the first derivative evaluation: the
// entries go to memory for the later sweeps and, still in registers, through the same
// arithmetic a sweep applies to them after reading them back
template <int C>
__device__ __forceinline__ void blo_sumtable_sweep_tile(const BloView &v, const ModelConst &m,
                                                        unsigned int pclv, unsigned int cclv,
                                                        double *sumtable, unsigned int tile,
                                                        const double (&dg)[4][3], double out[3],
                                                        unsigned long long vec_end)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  double ck[kPasses][3];
#pragma unroll 1
  for (int half = 0; half < kPasses; half += 2)
  {
    double p[2][4], c[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = min(site0 + (half + j) * kSitesPerPass, v.s.sites - 1);
      blo_load_entry<C>(v, pclv, site, cat, p[j]);
      blo_load_entry<C>(v, cclv, site, cat, c[j]);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = site0 + (half + j) * kSitesPerPass;
      double sv[4];
      sumtable_entry<C>(m, p[j], c[j], sv);
      if (site < v.s.sites) st256(sumtable + ((size_t)site * C + cat) * 4, sv);
      // the entry's terms of the first derivative evaluation, while it is in registers
      if (half == 0)
        sweep_entry_terms<C, false>(m, sv, dg, ck[j]);
      else
        sweep_entry_terms<C, false>(m, sv, dg, ck[2 + j]);
    }
  }
  unsigned int scv[kPasses];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps) scv[ps] = 0u;
  LnlProd lp;
  sweep_tile_finish<C, false>(m, ck, scv, v.s.sites, tile, out, lp, vec_end);
}

// edge log-likelihood of one tile (as edge_lnl_kernel); returns this thread's share
template <int C>
__device__ __forceinline__ double blo_lnl_tile(const BloView &v, const ModelConst &m, const BloEdge &e,
                                               unsigned int tile, const double *smat)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  const double *pm = smat + cat * kMatPad;
  double acc = 0.0;
#pragma unroll 1
  for (int half = 0; half < kPasses; half += 2)
  {
    double p[2][4], c[2][4];
    unsigned int sc[2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = min(site0 + (half + j) * kSitesPerPass, v.s.sites - 1);
      blo_load_entry<C>(v, e.p_clv, site, cat, p[j]);
      blo_load_entry<C>(v, e.c_clv, site, cat, c[j]);
      sc[j] = blo_load_scaler(v, e.p_scal, site) + blo_load_scaler(v, e.c_scal, site);
    }
    double f[2][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
      double row[4];
      lds256(pm + 4 * i, row);
#pragma unroll
      for (int j = 0; j < 2; ++j) f[j][i] = __dmul_rn(__dmul_rn(m.freqs[i], dot4(row, c[j])), p[j][i]);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int site = site0 + (half + j) * kSitesPerPass;
      const double tk = __dadd_rn(__dadd_rn(f[j][0], f[j][1]), __dadd_rn(f[j][2], f[j][3]));
      const double term = (tk > 0.0) ? __dmul_rn(tk, m.rate_weight) : 0.0;
      double terma = term;
#pragma unroll
      for (int k = 1; k < C; ++k)
        terma = __dadd_rn(terma, __shfl_down_sync(0xffffffffu, term, k, C));
      if (cat == 0 && site < v.s.sites)
      {
        double lk = log(terma);
        if (sc[j]) lk = __dadd_rn(lk, __dmul_rn((double)sc[j], m.log_scale_thr));
        if (m.weights) lk = __dmul_rn(lk, (double)__ldg(m.weights + site));
        acc = __dadd_rn(acc, lk);
      }
    }
  }
  return acc;
}

// ---- the persistent kernel ----------------------------------------------------
// CLUSTER = false: cooperative launch, phases separated by grid barriers.
// CLUSTER = true : the whole grid is ONE thread-block cluster (<= 8 CTAs) and phases are
// separated by the hardware cluster barrier (release/acquire at cluster scope, ~0.2 us
// instead of a ~2 us round of global atomics): for programs whose (candidate, tile) work
// fits a handful of CTAs -- OptimizeAllBranches on an alignment of a few thousand sites is
// several thousand barriers per call, and nothing else.
#ifndef PT_BLO_MINB
#define PT_BLO_MINB 2
#endif
template <int C, bool CLUSTER>
__global__ void __launch_bounds__(kSitesPerPass *C, (C <= 4 && !CLUSTER ? PT_BLO_MINB : 1)) blo_kernel(BloArgs a)
{
  constexpr int T = kSitesPerPass * C;
  extern __shared__ __align__(16) double ring[];
  __shared__ double red[2 * (T / 32 < 1 ? 1 : T / 32)];   // two sums per reduction (block_sum_n)
  __shared__ __align__(16) double smat[2 * C * kMatPad];
  __shared__ double sh_len;
  __shared__ int sh_done, sh_status;
  // fin[e] >= epoch <=> candidate e's solve of the current OPT step has finished.  The mark
  // only ever grows (one epoch per executed OPT step), so a CTA still testing "anybody
  // running?" for step k cannot be confused by a faster CTA arming step k+1.
  int epoch = 0;
  const int cat = threadIdx.x % C;
  struct Barrier
  {
    __device__ __forceinline__ void sync() const
    {
      if (CLUSTER)
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
      else
        cg::this_grid().sync();
    }
  } grid;

  // exchanges are numbered; the count is uniform over the grid (one per Newton phase and
  // per lnL phase, whether or not a CTA takes part) and carried from launch to launch
  const bool sharded = a.comm.nranks > 1;
  unsigned long long xq = sharded ? *(volatile unsigned long long *)a.comm.seq : 0ull;

  NewtonArgs na{};
  na.m = a.m;
  na.sumtables = a.sumtables;
  na.entry_stride = a.entry_stride;
  na.sclr = nullptr;
  na.st = a.nst;
  na.partials = a.partials;
  na.sites = a.sites; na.B = a.B; na.ntiles = a.ntiles; na.chunk_tiles = a.chunk_tiles;
  na.nchunks = a.nchunks; na.max_iter = a.max_newton;
  na.eps = a.xtol; na.lambda = a.lambda; na.want_lnl = 0; na.mode = a.procedure == 0 ? 1 : 0;
  na.xmin = a.xmin; na.xmax = a.xmax; na.lnl_margin = 0.0; na.vec_end = a.vec_end;

  auto live = [&](unsigned int e) {
    return __ldcg(&a.cand[e].active) && !__ldcg(&a.cand[e].failed);
  };

  // edge log-likelihood of every candidate for which want(e) holds -> store(e, value)
  auto lnl_phase = [&](const BloEdge *edges, auto want, auto store) {
    for (unsigned int w = blockIdx.x; w < a.B * a.ntiles; w += gridDim.x)
    {
      const unsigned int e = w / a.ntiles, tile = w % a.ntiles;
      if (!want(e)) continue;
      const BloEdge edge = edges[e];
      blo_stage_matrices<C>(a.v, edge.mat, edge.mat, 1, smat);
      const double part = block_sum<T>(blo_lnl_tile<C>(a.v, a.m, edge, tile, smat), red);
      if (threadIdx.x == 0) a.tile_partials[(size_t)e * a.ntiles + tile] = part;
    }
    grid.sync();
    for (unsigned int e = blockIdx.x; e < a.B; e += gridDim.x)
    {
      if (!want(e)) continue;
      double acc = 0.0;
      for (unsigned int t = threadIdx.x; t < a.ntiles; t += T)
        acc = __dadd_rn(acc, __ldcg(a.tile_partials + (size_t)e * a.ntiles + t));
      acc = block_sum<T>(acc, red);
      if (sharded)
      {
        // this shard's sum -> the sum over all shards (pll_optimize.c:1211-1212)
        if (threadIdx.x == 0) shard_put(a.comm, xq, e, acc);
        if (threadIdx.x < 32)
        {
          double s[1];
          shard_collect<1>(a.comm, xq, e, s);
          acc = s[0];
        }
      }
      if (threadIdx.x == 0)
      {
        // MAP: log(lambda) - lambda t per branch, added one by one in LogLikelihood's order
        // (PTPLL/pll_partition.cpp:318-341)
        for (unsigned int k = 0; k < a.n_prior; ++k)
          acc = __dadd_rn(acc, __dadd_rn(log(a.lambda),
                                         -__dmul_rn(a.lambda, __ldcg(a.lens + (size_t)e * a.NL + a.prior_slots[k]))));
        store(e, acc);
      }
    }
    if (sharded) xq += 1;
    grid.sync();
  };

  // CLV updates [k, k2) of candidate e on one tile, back to back.  No grid barrier is needed
  // between them: an entry written by one update is re-read by the next only by the thread
  // that wrote it (same tile, same thread <-> (site, category) mapping), the scaler word of a
  // site by the lanes next to its writer (ordered by the CTA barrier of the matrix staging).
  auto update_run_tile = [&](unsigned int e, unsigned int tile, unsigned int k, unsigned int k2) {
    for (unsigned int kk = k; kk < k2; ++kk)
    {
      const BloStep st = a.steps[(size_t)kk * a.B + e];
      if (st.type != BLO_OP) continue;
      blo_stage_matrices<C>(a.v, st.b_mat, st.c_mat, 2, smat);
      blo_op_tile<C>(a.v, st, tile, smat);
    }
  };

  auto wants = [&](unsigned int e, const BloStep &st) { return st.type == BLO_OPT && live(e); };
  auto start_length = [&](unsigned int e, const BloStep &st) {
    double guess = __ldcg(a.lens + (size_t)e * a.NL + st.slot);
    if (a.procedure == 0 && (guess < a.xmin || guess > a.xmax))
      guess = 0.1;  // PLLMOD_OPT_DEFAULT_BRANCH_LEN
    return guess;
  };
  const size_t pstride = (size_t)a.B * a.nchunks * 3;
  // a run of CLV updates can share a phase with the edge optimisation that follows it only
  // if both use the same work items ((candidate, chunk): what a thread wrote is what it reads),
  // i.e. when a chunk is a tile (alignments up to 64k sites -- where barriers are what costs)
  const bool chunk_items_ok = a.chunk_tiles == 1;

  // Executes the steps from k on that form one group and returns the first step after it:
  //   MAT ... MAT          one phase
  //   OP ... OP [OPT]      the updates of a tile back to back; the optimisation's sumtable and
  //                        first derivative evaluation in the same phase where possible; then
  //                        one phase per further evaluation
  auto run_group = [&](unsigned int k, unsigned int kend) -> unsigned int {
    const int type = a.step_type[k];
    if (type == BLO_MAT)
    {
      // P matrices of edges at their current lengths (a traversal's pll_update_prob_matrices)
      unsigned int k2 = k;
      while (k2 < kend && a.step_type[k2] == BLO_MAT) ++k2;
      const unsigned int n = (k2 - k) * a.B * C;
      for (unsigned int w = blockIdx.x * T + threadIdx.x; w < n; w += gridDim.x * T)
      {
        const unsigned int c = w % C, e = (w / C) % a.B, kk = k + w / (C * a.B);
        const BloStep st = a.steps[(size_t)kk * a.B + e];
        if (st.type == BLO_MAT && live(e))
          pmatrix_block(a.m, __ldcg(a.lens + (size_t)e * a.NL + st.slot), c,
                        blo_mat_ptr<C>(a.v, st.b_mat) + c * 16);
      }
      grid.sync();
      return k2;
    }
    if (type != BLO_OP && type != BLO_OPT) return k + 1;

    unsigned int k2 = k;
    while (k2 < kend && a.step_type[k2] == BLO_OP) ++k2;
    const bool has_opt = k2 < kend && a.step_type[k2] == BLO_OPT;
    const bool has_ops = k2 > k;
    const bool fused = has_opt && has_ops && chunk_items_ok;
    const BloStep *row = a.steps + (size_t)(has_opt ? k2 : k) * a.B;   // the OPT's row

    // Phase A (tile items): the updates alone, when they cannot share the optimisation's work
    // items.  Phase B (chunk items): the optimisation's phase 0 -- arm the solve, build the
    // sumtable and evaluate the derivatives at the start length from the entries while they
    // are in registers -- preceded, tile by tile, by the updates if they are fused in.
    // (One loop body for both so that the tile code exists once.)
    for (int phase = 0; phase < 2; ++phase)
    {
      const bool by_chunk = phase == 1;
      const bool ops_here = by_chunk ? fused : (has_ops && !fused);
      const bool opt_here = by_chunk && has_opt;
      if (!ops_here && !opt_here) continue;
      if (opt_here)
      {
        epoch += 1;
        for (unsigned int e = blockIdx.x; e < a.B; e += gridDim.x)
        {
          if (threadIdx.x != 0) continue;
          NewtonState ns{};
          const BloStep st = row[e];
          const bool on = wants(e, st);
          if (on) ns.len = start_length(e, st); else ns.done = 1;
          a.nst[e] = ns;
          if (!on) a.fin[e] = epoch;
        }
      }
      const unsigned int per = by_chunk ? a.nchunks : a.ntiles;
      for (unsigned int w = blockIdx.x; w < a.B * per; w += gridDim.x)
      {
        const unsigned int e = w / per, item = w % per;
        if (!live(e)) continue;
        const BloStep st = row[e];
        const bool on = opt_here && st.type == BLO_OPT;
        if (!on && !ops_here) continue;
        const unsigned int t0 = by_chunk ? item * a.chunk_tiles : item;
        const unsigned int t1 = by_chunk ? min(t0 + a.chunk_tiles, a.ntiles) : item + 1;
        if (ops_here)
          for (unsigned int tile = t0; tile < t1; ++tile) update_run_tile(e, tile, k, k2);
        if (!on) continue;
        double dg[4][3];
        diag_regs<C>(a.m, start_length(e, st), cat, dg);
        double out[3] = {0.0, 0.0, 0.0};
        for (unsigned int tile = t0; tile < t1; ++tile)
          blo_sumtable_sweep_tile<C>(a.v, a.m, st.a_clv, st.b_clv, a.sumtables + (size_t)e * a.entry_stride,
                                     tile, dg, out, a.vec_end);
        double s12[2] = {out[0], out[1]};
        block_sum_n<T, 2>(s12, red);   // the same tree per value as block_sum: same bits, half the barriers
        if (threadIdx.x == 0)
        {
          a.partials[(size_t)w * 3 + 0] = s12[0];
          a.partials[(size_t)w * 3 + 1] = s12[1];
        }
      }
      grid.sync();
    }
    if (!has_opt) return k2;

    // Edge optimisation, one grid barrier per further derivative evaluation.  Phase p >= 1:
    // EVERY CTA that holds a chunk of candidate e first redoes e's reduction and scalar
    // Newton step from evaluation p-1 (same inputs, same code => same result in all of
    // them; only the CTA of chunk 0 publishes it), then sweeps its chunk at the new
    // length.  Partials and Newton state alternate between two buffers so that a fast CTA
    // never overwrites what a slow one still reads.
    for (unsigned int p = 1; p <= a.max_newton; ++p)
    {
      const double *prd = a.partials + ((p - 1) & 1) * pstride;
      double *pwr = a.partials + (p & 1) * pstride;
      const NewtonState *srd = a.nst + ((p - 1) & 1) * (size_t)a.B;
      NewtonState *swr = a.nst + (p & 1) * (size_t)a.B;
      if (sharded)
      {
        // Site shards: the (d1, d2) sums of evaluation p-1 over THIS shard's chunks go out to
        // all shards (pll_optimize.c:1249-1255), one publishing CTA per candidate, before any
        // CTA of this grid starts to wait for a sum -- publishing never waits, so the order
        // in which CTAs reach their work items cannot close a cycle.
        for (unsigned int e = blockIdx.x; e < a.B; e += gridDim.x)
        {
          if (__syncthreads_or(threadIdx.x == 0 && __ldcg(a.fin + e) >= epoch)) continue;
          double a1 = 0.0, a2 = 0.0;
          for (unsigned int t = threadIdx.x; t < a.nchunks; t += T)
          {
            a1 = __dadd_rn(a1, __ldcg(prd + ((size_t)e * a.nchunks + t) * 3 + 0));
            a2 = __dadd_rn(a2, __ldcg(prd + ((size_t)e * a.nchunks + t) * 3 + 1));
          }
          double a12[2] = {a1, a2};
          block_sum_n<T, 2>(a12, red);
          a1 = a12[0];
          a2 = a12[1];
          if (threadIdx.x == 0)
          {
            shard_put(a.comm, xq, 2 * e, a1);
            shard_put(a.comm, xq, 2 * e + 1, a2);
          }
        }
      }
      for (unsigned int w = blockIdx.x; w < a.B * a.nchunks; w += gridDim.x)
      {
        const unsigned int e = w / a.nchunks, chunk = w % a.nchunks;
        // finished in an earlier phase -- or in this one, already published by the CTA of
        // chunk 0: either way there is nothing left to sweep
        // (the mark may be written while we look: one thread reads, the CTA follows it)
        if (__syncthreads_or(threadIdx.x == 0 && __ldcg(a.fin + e) >= epoch)) continue;
        double a1 = 0.0, a2 = 0.0;
        if (sharded)
        {
          // the sums over all shards, in rank order (same bits on every shard and in every CTA)
          if (threadIdx.x < 32)
          {
            double s[2];
            shard_collect<2>(a.comm, xq, 2 * e, s);
            a1 = s[0];
            a2 = s[1];
          }
        }
        else
        {
          for (unsigned int t = threadIdx.x; t < a.nchunks; t += T)
          {
            a1 = __dadd_rn(a1, __ldcg(prd + ((size_t)e * a.nchunks + t) * 3 + 0));
            a2 = __dadd_rn(a2, __ldcg(prd + ((size_t)e * a.nchunks + t) * 3 + 1));
          }
          double a12[2] = {a1, a2};
          block_sum_n<T, 2>(a12, red);
          a1 = a12[0];
          a2 = a12[1];
        }
        NewtonState ns;
        if (threadIdx.x == 0)
        {
          ns = newton_scalar_step(na, srd + e, a1, a2, 0.0);
          sh_len = ns.len;
          sh_done = ns.done;
          sh_status = ns.status;
          if (chunk == 0) swr[e] = ns;
        }
        __syncthreads();
        const double len = sh_len;
        const int done = sh_done;
        if (done)
        {
          if (chunk == 0)
          {
            // accept the result (pll_optimize.c:824-868 / PTPLL/pll_partition.cpp:483-487):
            // length slot + P matrix, or the failure (pll_errno / std::range_error)
            const BloStep st = row[e];
            const double old = __ldcg(a.lens + (size_t)e * a.NL + st.slot);
            const int status = sh_status;
            __syncthreads();  // everybody has read the old length before thread 0 replaces it
            // procedure 1: UpdateBranchLength always recomputes the matrix (PTPLL/pll_partition.cpp:395-406)
            if (status == 0 && (a.procedure == 1 || fabs(old - len) > 1e-10) && threadIdx.x < C)
              pmatrix_block(a.m, len, threadIdx.x, blo_mat_ptr<C>(a.v, st.b_mat) + threadIdx.x * 16);
            if (threadIdx.x == 0)
            {
              if (status != 0)
                a.cand[e].failed = status;
              else
                a.lens[(size_t)e * a.NL + st.slot] = len;
              a.cand[e].newton_evals += ns.iters;
              a.fin[e] = epoch;
            }
          }
          __syncthreads();   // sh_len / sh_done are rewritten by the next work item
          continue;
        }
        double dg[4][3];
        diag_regs<C>(a.m, len, cat, dg);
        double out[3] = {0.0, 0.0, 0.0};
        LnlProd lp;
        const unsigned int t1 = min((chunk + 1) * a.chunk_tiles, a.ntiles);
        sweep_chunk<C, false>(a.m, a.sumtables + (size_t)e * a.entry_stride, nullptr, a.sites,
                              chunk * a.chunk_tiles, t1, dg, ring, out, lp, a.vec_end);
        double s12[2] = {out[0], out[1]};
        block_sum_n<T, 2>(s12, red);
        if (threadIdx.x == 0)
        {
          pwr[(size_t)w * 3 + 0] = s12[0];
          pwr[(size_t)w * 3 + 1] = s12[1];
        }
      }
      if (sharded) xq += 1;
      grid.sync();
      int running = 0;
      for (unsigned int e = threadIdx.x; e < a.B; e += T) running |= __ldcg(a.fin + e) < epoch;
      if (!__syncthreads_or(running)) break;
    }
    return k2 + 1;
  };

  for (unsigned int e = blockIdx.x * T + threadIdx.x; e < a.B; e += gridDim.x * T) a.fin[e] = 0;
  // private P matrices at the start lengths (in-place programs use the engine's matrices as
  // they are: a stale one stays stale until a step recomputes it, as in the call sequence)
  for (unsigned int w = blockIdx.x * T + threadIdx.x; a.v.pmat && w < a.B * a.NL * C; w += gridDim.x * T)
  {
    const unsigned int c = w % C, slot = (w / C) % a.NL, e = w / (C * a.NL);
    pmatrix_block(a.m, a.lens[(size_t)e * a.NL + slot], c,
                  a.v.pmat + ((size_t)(e * a.NL + slot) * C + c) * 16);
  }
  grid.sync();
  // round -1: the prologue and the lnL before the first smoothing (pll_optimize.c:1000-1007);
  // rounds 0..: one smoothing / pass each, its lnL and the stop decision; last: the final lnL
  // (Partition::LogLikelihood(node)) on whatever state each candidate ended in.  One loop so
  // that run_step and lnl_phase have a single call site each (they inline; BloArgs stays in
  // the parameter bank instead of being copied to the stack).
  bool finishing = false;
  for (int round = -1;; ++round)
  {
    if (!finishing)
    {
      // local optimisation: the first smoothing still reads CLVs the engine holds, from the
      // second on every CLV the body has touched is the candidate's private copy; whole-tree
      // optimisation: the traversals of a pass depend on the orientation the previous one left
      const unsigned int variant = round < 0 ? 0u : min((unsigned int)round, a.n_bodies - 1);
      const unsigned int k0 = round < 0 ? 0u : a.n_pro + variant * a.n_body;
      const unsigned int k1 = k0 + (round < 0 ? a.n_pro : a.n_body);
      for (unsigned int k = k0; k < k1;) k = run_group(k, k1);
    }
    lnl_phase(finishing ? a.final_edge : a.smooth_edge,
              [&](unsigned int e) { return round < 0 || finishing || live(e); },
              [&](unsigned int e, double now) {
                if (finishing) { a.cand[e].final_lnl = now; return; }
                if (round < 0) { a.cand[e].lnl = now; return; }
                BloCand c = a.cand[e];
                if (a.procedure == 0)
                {
                  // pll_optimize.c:1058-1076
                  if (!(now - c.lnl > now * 1e-14))
                    c.active = 0;
                  else
                  {
                    c.smoothings += 1;
                    if (fabs(now - c.lnl) < a.tolerance || (unsigned int)c.smoothings >= a.smoothings) c.active = 0;
                  }
                }
                else
                {
                  // PTPLL/pll_partition.cpp:566-581: another pass while |before - after| > EPSILON
                  c.smoothings += 1;
                  if (!(fabs(c.lnl - now) > a.tolerance) || (unsigned int)c.smoothings >= a.smoothings) c.active = 0;
                }
                c.lnl = now;
                a.cand[e] = c;
              });
    if (finishing) break;
    if (round >= 0)
    {
      int running = 0;
      for (unsigned int e = threadIdx.x; e < a.B; e += T) running |= live(e);
      if (!__syncthreads_or(running)) finishing = true;
    }
    if (round + 1 >= (int)a.smoothings) finishing = true;
  }
  // every CTA has read the sequence number long ago (grid barriers since); the next launch on
  // this stream continues the numbering
  if (sharded && blockIdx.x == 0 && threadIdx.x == 0) *a.comm.seq = xq;
}

} // namespace ptgpu
