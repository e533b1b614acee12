/*
 * oracle/pt_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, single-threaded CPU restatement of the arithmetic the reference
 * (matsengrp/pt -> libptpll -> libpll, AVX code path) performs on pt's NNI
 * scoring hot path.  It is the checker the CUDA engine is compared against.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product (pt_b200/, include/) never
 * links, imports or calls anything in oracle/.
 *
 * PARITY PIN: this restatement is checked bit-for-bit against the reference's
 * own code compiled in place (oracle/_ref/libpllref.so, built by
 * oracle/Makefile.ref) by tests/golden/make_golden.py, and against the golden
 * vectors that script commits under tests/golden/ (tests/test_oracle.py).
 *
 * All paths in comments are under /root/reference/; LIBPLL =
 * lib/libptpll/lib/pll-modules/libs/libpll/src, PTPLL = lib/libptpll/src.
 * Layouts (LIBPLL/pll.c:527-573,730-805):
 *   clv[site][cat][state]  4 states, C rate categories
 *   pmatrix[cat][i][j]     row i = parent state
 *   sumtable[site][cat][j]
 *   scaler[site]           unsigned counts of 2^256 rescalings
 * Build with -ffp-contract=off: the reference's AVX path never fuses a
 * multiply with an add (compiled with -mavx only, LIBPLL/Makefile.am:52-64).
 */
#ifndef PT_ORACLE_H
#define PT_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* -- model (LIBPLL/models.c:24-331, LIBPLL/gamma.c:24-292) -- */
void pto_eigen_gtr(const double freqs[4], const double subst[6],
                   double evecs[16], double inv_evecs[16], double evals[4]);
int pto_gamma_rates_mean(double alpha, unsigned int cats, double *rates);

/* -- tips (LIBPLL/pll.c:905-939, LIBPLL/maps.c:46-64) -- */
unsigned int pto_nt_mask(int ch); /* 0 = illegal */
int pto_tip_clv(const char *seq, unsigned int sites, unsigned int cats,
                double *clv);

/* -- P matrix (LIBPLL/core_pmatrix_avx.c:42-310) -- */
void pto_pmatrix(const double evals[4], const double evecs[16],
                 const double inv_evecs[16], const double *rates,
                 unsigned int cats, double t, double *pmat);

/* -- CLV update (LIBPLL/core_partials_avx.c:24-46,366-529) --
   any scaler pointer may be NULL */
void pto_update_partial(unsigned int sites, unsigned int cats,
                        double *parent_clv, unsigned int *parent_scaler,
                        const double *left_clv, const double *right_clv,
                        const double *lmat, const double *rmat,
                        const unsigned int *left_scaler,
                        const unsigned int *right_scaler);

/* -- edge log-likelihood (LIBPLL/core_likelihood_avx.c:1080-1266) --
   rate weights are 1/cats, pattern weights 1 (pt never compresses) */
double pto_edge_lnl(unsigned int sites, unsigned int cats,
                    const double *clvp, const unsigned int *sp,
                    const double *clvc, const unsigned int *sc,
                    const double *pmat, const double freqs[4],
                    double *persite);

/* -- sumtable (LIBPLL/core_derivatives_avx.c:25-207) -- */
void pto_sumtable(unsigned int sites, unsigned int cats,
                  const double *clvp, const double *clvc,
                  const double evecs[16], const double inv_evecs[16],
                  const double freqs[4], double *sumtable);

/* -- derivatives of -lnL (LIBPLL/core_derivatives.c:560-575,
      LIBPLL/core_derivatives_avx.c:898-1174) -- */
void pto_derivatives(unsigned int sites, unsigned int cats,
                     const double *sumtable, const double evals[4],
                     const double *rates, double t, double *d1, double *d2);

/* -- pt's own Newton loop (PTPLL/pll_partition.cpp:408-488) --
   returns the new length; *iters = derivative evaluations done;
   *status: 0 ok, 1 d1 not finite, 2 d2 not finite (reference throws) */
double pto_optimize_branch(unsigned int sites, unsigned int cats,
                           const double *sumtable, const double evals[4],
                           const double *rates, double len0,
                           unsigned int max_iter, double eps, double map_lambda,
                           int *iters, int *status);

/* -- the same three with pattern weights (LIBPLL/core_likelihood_avx.c:1254,
      LIBPLL/core_derivatives_avx.c:1123-1139,1153-1154); weights == NULL is all ones -- */
double pto_edge_lnl_w(unsigned int sites, unsigned int cats,
                      const double *clvp, const unsigned int *sp,
                      const double *clvc, const unsigned int *sc,
                      const double *pmat, const double freqs[4],
                      const unsigned int *weights, double *persite);
void pto_derivatives_w(unsigned int sites, unsigned int cats,
                       const double *sumtable, const double evals[4],
                       const double *rates, const unsigned int *weights, double t,
                       double *d1, double *d2);
double pto_optimize_branch_w(unsigned int sites, unsigned int cats,
                             const double *sumtable, const double evals[4],
                             const double *rates, const unsigned int *weights, double len0,
                             unsigned int max_iter, double eps, double map_lambda,
                             int *iters, int *status);

/* -- pll_compress_site_patterns for nucleotide data (LIBPLL/compress.c:138-282) --
   sequences are overwritten by the unique columns (NUL-terminated), *length by their
   number, weight[0 .. *length) by their multiplicities; -1 on an illegal character */
int pto_compress_site_patterns(char **sequence, unsigned int count, unsigned int *length,
                               unsigned int *weight);

#ifdef __cplusplus
}
#endif
#endif
