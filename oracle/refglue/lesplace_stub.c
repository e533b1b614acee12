/*
 * oracle/refglue/lesplace_stub.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * lesplace needs GSL, which this image does not have.  Marginal-likelihood
 * mode (--marginal) is off in every BASELINE config and its only pt test is
 * compiled out (test/test_ptw.cpp:506-521), so the entry point the reference
 * wrapper links against (PTPLL/pll_partition.cpp:302) is a loud stub.
 */
#include <stdio.h>
#include <stdlib.h>
#include <stddef.h>

double lesplace_gamma(double (*logP)(void *, size_t, double),
                      void (*dlogPs)(void *, size_t, double *, double *),
                      void *data, size_t count, double map, const double *mapes)
{
  (void)logP; (void)dlogPs; (void)data; (void)count; (void)map; (void)mapes;
  fprintf(stderr, "lesplace_gamma: unavailable in the oracle build (no GSL)\n");
  abort();
}

/* pll-modules' consensus code calls a second bison parser (PLLMOD/tree/
   split_utree.y) that cannot be generated here and is never reached from pt. */
void *pll_utree_split_newick_string(char *s, unsigned int tip_count, void *names_hash)
{
  (void)s; (void)tip_count; (void)names_hash;
  fprintf(stderr, "pll_utree_split_newick_string: unavailable in the oracle build\n");
  abort();
}
