/*
 * pt_b200/host/src/pll_errno.c -- libpll's thread-local error state for the host build.
 *
 * libpll defines its thread-local error state in pll.c next to
 * pll_partition_create and the rest of the CPU likelihood set-up
 * (LIBPLL/pll.c:24-25).  The B200 host build links only the reference's
 * host-side tree / FASTA objects, which still refer to these two variables, so
 * they are defined here instead of pulling pll.o (and with it the CPU kernels)
 * onto the link line.
 */
#include <libpll/pll.h>

__thread int pll_errno;
__thread char pll_errmsg[200] = {0};
