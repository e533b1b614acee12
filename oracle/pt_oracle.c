/*
 * oracle/pt_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * See pt_oracle.h for scope, layouts, the parity pin and citation prefixes.
 * Compile:  gcc -O2 -ffp-contract=off -fPIC -shared -o libpt_oracle.so pt_oracle.c -lm
 */
#include "pt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NS 4 /* states */

/* 2^256 and 2^-256: LIBPLL/pll.h PLL_SCALE_FACTOR / PLL_SCALE_THRESHOLD */
static double scale_factor(void) { return ldexp(1.0, 256); }
static double scale_threshold(void) { return 1.0 / ldexp(1.0, 256); }

/* (p0 + p1) + (p2 + p3) with separately rounded products: the summation
   tree of every 4-term dot product on the reference's AVX path (SURVEY N1;
   unpack/add/permute/blend/add, e.g. LIBPLL/core_partials_avx.c:460-471) */
static inline double dot4(const double *a, const double *b)
{
  double p0 = a[0] * b[0], p1 = a[1] * b[1];
  double p2 = a[2] * b[2], p3 = a[3] * b[3];
  return (p0 + p1) + (p2 + p3);
}

/* ------------------------------------------------------------------ */
/* model: symmetric eigensystem of sqrt(pi) Q sqrt(pi)^-1             */
/* ------------------------------------------------------------------ */

#define A(r, c) a[(r) * NS + (c)]

/* Householder tridiagonalisation, column-oriented variant the reference
   uses (LIBPLL/models.c:100-178), indices shifted to 0-based. */
static void householder_tridiag(double *a, double *d, double *e)
{
  const int n = NS;
  for (int i = n - 1; i >= 1; --i)
  {
    const int l = i; /* rows 0..l-1 of column i are reduced */
    double h = 0.0, scale = 0.0;
    if (l > 1)
    {
      for (int k = 0; k < l; ++k) scale += fabs(A(k, i));
      if (scale == 0.0)
        e[i] = A(l - 1, i);
      else
      {
        for (int k = 0; k < l; ++k)
        {
          A(k, i) /= scale;
          h += A(k, i) * A(k, i);
        }
        double f = A(l - 1, i);
        double g = (f > 0) ? -sqrt(h) : sqrt(h);
        e[i] = scale * g;
        h -= f * g;
        A(l - 1, i) = f - g;
        f = 0.0;
        for (int j = 0; j < l; ++j)
        {
          A(i, j) = A(j, i) / h;
          g = 0.0;
          for (int k = 0; k <= j; ++k) g += A(k, j) * A(k, i);
          for (int k = j + 1; k < l; ++k) g += A(j, k) * A(k, i);
          e[j] = g / h;
          f += e[j] * A(j, i);
        }
        const double hh = f / (h + h);
        for (int j = 0; j < l; ++j)
        {
          f = A(j, i);
          g = e[j] - hh * f;
          e[j] = g;
          for (int k = 0; k <= j; ++k)
            A(k, j) -= (f * e[k] + g * A(k, i));
        }
      }
    }
    else
      e[i] = A(l - 1, i);
    d[i] = h;
  }
  d[0] = 0.0;
  e[0] = 0.0;
  for (int i = 0; i < n; ++i)
  {
    const int l = i;
    if (d[i] != 0.0)
    {
      for (int j = 0; j < l; ++j)
      {
        double g = 0.0;
        for (int k = 0; k < l; ++k) g += A(k, i) * A(j, k);
        for (int k = 0; k < l; ++k) A(j, k) -= g * A(i, k);
      }
    }
    d[i] = A(i, i);
    A(i, i) = 1.0;
    for (int j = 0; j < l; ++j) A(i, j) = A(j, i) = 0.0;
  }
}

/* implicit-shift QL on the tridiagonal (d, e), rotating the rows of a
   (LIBPLL/models.c:24-97), 0-based */
static void ql_implicit(double *d, double *e, double *a)
{
  const int n = NS;
  for (int i = 1; i < n; ++i) e[i - 1] = e[i];
  e[n - 1] = 0.0;

  for (int l = 0; l < n; ++l)
  {
    int m;
    do
    {
      for (m = l; m < n - 1; ++m)
      {
        const double dd = fabs(d[m]) + fabs(d[m + 1]);
        if (fabs(e[m]) + dd == dd) break;
      }
      if (m != l)
      {
        double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
        double r = sqrt((g * g) + 1.0);
        g = d[m] - d[l] + e[l] / (g + ((g < 0) ? -fabs(r) : fabs(r)));
        double s = 1.0, c = 1.0, p = 0.0;
        for (int i = m - 1; i >= l; --i)
        {
          double f = s * e[i];
          const double b = c * e[i];
          if (fabs(f) >= fabs(g))
          {
            c = g / f;
            r = sqrt((c * c) + 1.0);
            e[i + 1] = f * r;
            s = 1.0 / r;
            c *= s;
          }
          else
          {
            s = f / g;
            r = sqrt((s * s) + 1.0);
            e[i + 1] = g * r;
            c = 1.0 / r;
            s *= c;
          }
          g = d[i + 1] - p;
          r = (d[i] - g) * s + 2.0 * c * b;
          p = s * r;
          d[i + 1] = g + p;
          g = c * r - b;
          for (int k = 0; k < n; ++k)
          {
            f = A(i + 1, k);
            A(i + 1, k) = s * A(i, k) + c * f;
            A(i, k) = c * A(i, k) - s * f;
          }
        }
        d[l] = d[l] - p;
        e[l] = g;
        e[m] = 0.0;
      }
    } while (m != l);
  }
}

void pto_eigen_gtr(const double freqs[4], const double subst[6],
                   double evecs[16], double inv_evecs[16], double evals[4])
{
  /* LIBPLL/models.c:182-249: rates normalised by the last one, symmetrised
     rate matrix, mean substitution rate scaled to 1 */
  double rate[6], a[16], d[4], e[4];
  memcpy(rate, subst, sizeof rate);
  if (rate[5] > 0.0)
    for (int i = 0; i < 6; ++i) rate[i] /= rate[5];

  for (int i = 0; i < 16; ++i) a[i] = 0.0;
  int k = 0;
  for (int i = 0; i < NS; ++i)
    for (int j = i + 1; j < NS; ++j)
    {
      const double factor = rate[k++];
      A(i, j) = A(j, i) = factor * sqrt(freqs[i] * freqs[j]);
      A(i, i) -= factor * freqs[j];
      A(j, j) -= factor * freqs[i];
    }
  double mean = 0;
  for (int i = 0; i < NS; ++i) mean += freqs[i] * (-A(i, i));
  for (int i = 0; i < 16; ++i) a[i] /= mean;

  householder_tridiag(a, d, e);
  ql_implicit(d, e, a);

  /* LIBPLL/models.c:290-318 */
  memcpy(evecs, a, sizeof(double) * 16);
  memcpy(evals, d, sizeof(double) * 4);
  for (int i = 0; i < NS; ++i)
    for (int j = 0; j < NS; ++j) inv_evecs[i * NS + j] = evecs[j * NS + i];
  for (int i = 0; i < NS; ++i)
    for (int j = 0; j < NS; ++j) inv_evecs[i * NS + j] /= sqrt(freqs[i]);
  for (int i = 0; i < NS; ++i)
    for (int j = 0; j < NS; ++j) evecs[i * NS + j] *= sqrt(freqs[j]);
}
#undef A

/* ------------------------------------------------------------------ */
/* discrete Gamma mean rates (LIBPLL/gamma.c:24-292; Yang 1994,        */
/* AS32 / AS91 / AS70 / ACM291)                                        */
/* ------------------------------------------------------------------ */

static double ln_gamma_fn(double alpha)
{
  /* LIBPLL/gamma.c:91-124 */
  double x = alpha, f = 0.0, z;
  if (x < 7.0)
  {
    f = 1.0;
    z = alpha - 1.0;
    while ((z = z + 1.0) < 7.0) f *= z;
    x = z;
    f = -log(f);
  }
  z = 1 / (x * x);
  return f + (x - 0.5) * log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z +
          .083333333333333) / x;
}

static double incomplete_gamma(double x, double alpha, double ln_gamma_alpha)
{
  /* LIBPLL/gamma.c:27-90 */
  const double p = alpha, g = ln_gamma_alpha;
  const double accurate = 1e-8, overflow = 1e30;
  double pn[6];

  if (x == 0) return 0;
  if (x < 0 || p <= 0) return -1;

  const double factor = exp(p * log(x) - x - g);
  if (!(x > 1 && x >= p))
  {
    /* series expansion */
    double gin = 1, term = 1, rn = p;
    do
    {
      rn++;
      term *= x / rn;
      gin += term;
    } while (term > accurate);
    gin *= factor / p;
    return gin;
  }
  /* continued fraction */
  double a = 1 - p, b = a + x + 1, term = 0;
  pn[0] = 1; pn[1] = x; pn[2] = x + 1; pn[3] = x * b;
  double gin = pn[2] / pn[3];
  for (;;)
  {
    a++;
    b += 2;
    term++;
    const double an = a * term;
    for (int i = 0; i < 2; ++i) pn[i + 4] = b * pn[i + 2] - an * pn[i];
    if (pn[5] != 0)
    {
      const double rn = pn[4] / pn[5];
      const double dif = fabs(gin - rn);
      if (!(dif > accurate) && dif <= accurate * rn)
        return 1 - factor * gin;
      gin = rn;
    }
    for (int i = 0; i < 4; ++i) pn[i] = pn[i + 2];
    if (!(fabs(pn[4]) < overflow))
      for (int i = 0; i < 4; ++i) pn[i] /= overflow;
  }
}

static double point_normal(double prob)
{
  /* LIBPLL/gamma.c:126-152 */
  const double a0 = -.322232431088, a1 = -1, a2 = -.342242088547,
               a3 = -.0204231210245, a4 = -.453642210148e-4,
               b0 = .0993484626060, b1 = .588581570495, b2 = .531103462366,
               b3 = .103537752850, b4 = .0038560700634;
  const double p = prob, p1 = (p < 0.5 ? p : 1 - p);
  if (p1 < 1e-20) return -9999;
  const double y = sqrt(log(1 / (p1 * p1)));
  const double z = y + ((((y * a4 + a3) * y + a2) * y + a1) * y + a0) /
                           ((((y * b4 + b3) * y + b2) * y + b1) * y + b0);
  return (p < 0.5 ? -z : z);
}

static double point_chi2(double prob, double v)
{
  /* LIBPLL/gamma.c:154-218 */
  const double e = .5e-6, aa = .6931471805, p = prob;
  double ch, a, q, p1, p2, t, x, b;

  if (p < .000002 || p > .999998 || v <= 0) return -1;
  const double g = ln_gamma_fn(v / 2);
  const double xx = v / 2, c = xx - 1;

  if (!(v >= -1.24 * log(p)))
  {
    ch = pow((p * xx * exp(g + xx * aa)), 1 / xx);
    if (ch - e < 0) return ch;
  }
  else if (v > .32)
  {
    x = point_normal(p);
    p1 = 0.222222 / v;
    ch = v * pow((x * sqrt(p1) + 1 - p1), 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (log(1 - p) - c * log(.5 * ch) + g);
  }
  else
  {
    ch = 0.4;
    a = log(1 - p);
    do
    {
      q = ch;
      p1 = 1 + ch * (4.67 + ch);
      p2 = ch * (6.73 + ch * (6.66 + ch));
      t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - exp(a + g + .5 * ch + c * aa) * p2 / p1) / t;
    } while (!(fabs(q / ch - 1) - .01 <= 0));
  }
  do
  {
    q = ch;
    p1 = .5 * ch;
    if ((t = incomplete_gamma(p1, xx, g)) < 0.0) return -1;
    p2 = p - t;
    t = p2 * exp(xx * aa + g + p1 - c * log(ch));
    b = t / ch;
    a = 0.5 * t - b * c;
    const double s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    const double s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    const double s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    const double s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    const double s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    const double s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 -
               b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (fabs(q / ch - 1) > e);
  return ch;
}

int pto_gamma_rates_mean(double alpha, unsigned int cats, double *rates)
{
  /* LIBPLL/gamma.c:220-292, PLL_GAMMA_RATES_MEAN branch */
  if (alpha < 0.02 || cats < 1) return 0;
  if (cats == 1) { rates[0] = 1.0; return 1; }
  const double factor = alpha / alpha * cats, alfa = alpha, beta = alpha;
  double *gp = (double *)malloc(cats * sizeof(double));
  const double lnga1 = ln_gamma_fn(alfa + 1);
  for (unsigned int i = 0; i < cats - 1; ++i)
    gp[i] = point_chi2((i + 1.0) / cats, 2.0 * (alfa)) / (2.0 * (beta));
  for (unsigned int i = 0; i < cats - 1; ++i)
    gp[i] = incomplete_gamma(gp[i] * beta, alfa + 1, lnga1);
  rates[0] = gp[0] * factor;
  rates[cats - 1] = (1 - gp[cats - 2]) * factor;
  for (unsigned int i = 1; i < cats - 1; ++i) rates[i] = (gp[i] - gp[i - 1]) * factor;
  free(gp);
  return 1;
}

/* ------------------------------------------------------------------ */
/* tips                                                                */
/* ------------------------------------------------------------------ */

unsigned int pto_nt_mask(int ch)
{
  /* LIBPLL/maps.c:46-64 (pll_map_nt): bit0=A bit1=C bit2=G bit3=T,
     IUPAC ambiguity codes as unions, gap/N/?/X/O = 15, case-insensitive */
  switch (ch >= 'a' && ch <= 'z' ? ch - 32 : ch)
  {
    case 'A': return 1;  case 'C': return 2;  case 'G': return 4;
    case 'T': case 'U': return 8;
    case 'M': return 3;  case 'R': return 5;  case 'S': return 6;
    case 'V': return 7;  case 'W': return 9;  case 'Y': return 10;
    case 'H': return 11; case 'K': return 12; case 'D': return 13;
    case 'B': return 14;
    case 'N': case 'O': case 'X': case '-': case '?': return 15;
    default: return 0;
  }
}

int pto_tip_clv(const char *seq, unsigned int sites, unsigned int cats, double *clv)
{
  /* LIBPLL/pll.c:905-939: 0/1 doubles, replicated for every rate category */
  for (unsigned int n = 0; n < sites; ++n)
  {
    unsigned int c = pto_nt_mask((unsigned char)seq[n]);
    if (!c) return 0;
    for (unsigned int k = 0; k < cats; ++k)
      for (unsigned int j = 0; j < NS; ++j)
        clv[(n * cats + k) * NS + j] = (double)((c >> j) & 1u);
  }
  return 1;
}

/* ------------------------------------------------------------------ */
/* P matrix                                                            */
/* ------------------------------------------------------------------ */

void pto_pmatrix(const double evals[4], const double evecs[16],
                 const double inv_evecs[16], const double *rates,
                 unsigned int cats, double t, double *pmat)
{
  /* LIBPLL/core_pmatrix_avx.c:42-310.  prop_invar is 0 in pt.
     t == 0 -> identity (:103-111).  Otherwise
       e_k   = expm1((lambda_k * r_c) * t)                 (:117-147)
       P_ij  = [(w_i0 v_0j + w_i1 v_1j) + (w_i2 v_2j + w_i3 v_3j)] + I_ij
       w_ik  = inv_evecs[i][k] * e_k                       (:173-183, :222) */
  for (unsigned int c = 0; c < cats; ++c, pmat += 16)
  {
    if (!t)
    {
      for (int i = 0; i < 16; ++i) pmat[i] = 0.0;
      pmat[0] = pmat[5] = pmat[10] = pmat[15] = 1;
      continue;
    }
    double e[4];
    for (int k = 0; k < NS; ++k) e[k] = expm1((evals[k] * rates[c]) * t);
    for (int i = 0; i < NS; ++i)
    {
      double w[4];
      for (int k = 0; k < NS; ++k) w[k] = inv_evecs[i * NS + k] * e[k];
      for (int j = 0; j < NS; ++j)
      {
        const double p0 = w[0] * evecs[0 * NS + j], p1 = w[1] * evecs[1 * NS + j];
        const double p2 = w[2] * evecs[2 * NS + j], p3 = w[3] * evecs[3 * NS + j];
        pmat[i * NS + j] = ((p0 + p1) + (p2 + p3)) + (i == j ? 1.0 : 0.0);
      }
    }
  }
}

/* ------------------------------------------------------------------ */
/* CLV update                                                          */
/* ------------------------------------------------------------------ */

void pto_update_partial(unsigned int sites, unsigned int cats,
                        double *parent_clv, unsigned int *parent_scaler,
                        const double *left_clv, const double *right_clv,
                        const double *lmat, const double *rmat,
                        const unsigned int *left_scaler,
                        const unsigned int *right_scaler)
{
  /* LIBPLL/core_partials_avx.c:366-529, per-site scaling mode.
     parent scaler starts as left + right (missing = 0) (:24-46,409);
     if every one of the 4*cats entries of a site is < 2^-256 (ordered
     compare, so NaN never scales) the site is multiplied by 2^256 and its
     count incremented (:491-527). */
  const double thr = scale_threshold(), fac = scale_factor();
  const unsigned int span = NS * cats;
  for (unsigned int n = 0; n < sites; ++n)
  {
    if (parent_scaler)
      parent_scaler[n] = (left_scaler ? left_scaler[n] : 0u) +
                         (right_scaler ? right_scaler[n] : 0u);
    int all_small = 1;
    for (unsigned int k = 0; k < cats; ++k)
    {
      const double *l = left_clv + (size_t)n * span + k * NS;
      const double *r = right_clv + (size_t)n * span + k * NS;
      double *p = parent_clv + (size_t)n * span + k * NS;
      for (int i = 0; i < NS; ++i)
      {
        const double x = dot4(lmat + k * 16 + i * NS, l);
        const double y = dot4(rmat + k * 16 + i * NS, r);
        p[i] = x * y;
        all_small &= (p[i] < thr);
      }
    }
    if (parent_scaler && all_small)
    {
      double *p = parent_clv + (size_t)n * span;
      for (unsigned int i = 0; i < span; ++i) p[i] *= fac;
      parent_scaler[n] += 1;
    }
  }
}

/* ------------------------------------------------------------------ */
/* edge log-likelihood                                                 */
/* ------------------------------------------------------------------ */

double pto_edge_lnl(unsigned int sites, unsigned int cats,
                    const double *clvp, const unsigned int *sp,
                    const double *clvc, const unsigned int *sc,
                    const double *pmat, const double freqs[4], double *persite)
{
  return pto_edge_lnl_w(sites, cats, clvp, sp, clvc, sc, pmat, freqs, NULL, persite);
}

double pto_edge_lnl_w(unsigned int sites, unsigned int cats,
                      const double *clvp, const unsigned int *sp,
                      const double *clvc, const unsigned int *sc,
                      const double *pmat, const double freqs[4],
                      const unsigned int *weights, double *persite)
{
  /* LIBPLL/core_likelihood_avx.c:1080-1266.
     t_k = (f0 + f1) + (f2 + f3), f_i = (pi_i * (P_k c)_i) * p_i  (:1194-1217)
     terma += t_k * w_k only when t_k > 0                          (:1225-1241)
     site = log(terma) + (s_p + s_c) * log(2^-256)                 (:1247-1251)
     logl accumulated in site order                                (:1259) */
  const double w = 1.0 / cats;
  const double log_thr = log(scale_threshold());
  double logl = 0;
  for (unsigned int n = 0; n < sites; ++n)
  {
    double terma = 0;
    for (unsigned int k = 0; k < cats; ++k)
    {
      const double *p = clvp + ((size_t)n * cats + k) * NS;
      const double *c = clvc + ((size_t)n * cats + k) * NS;
      double f[4];
      for (int i = 0; i < NS; ++i)
        f[i] = (freqs[i] * dot4(pmat + k * 16 + i * NS, c)) * p[i];
      const double tk = (f[0] + f[1]) + (f[2] + f[3]);
      if (tk > 0.) terma += tk * w;
    }
    double site = log(terma);
    const unsigned int sclr = (sp ? sp[n] : 0u) + (sc ? sc[n] : 0u);
    if (sclr) site += sclr * log_thr;
    site *= weights ? weights[n] : 1u; /* pattern weight (:1254) */
    if (persite) persite[n] = site;
    logl += site;
  }
  return logl;
}

/* ------------------------------------------------------------------ */
/* sumtable                                                            */
/* ------------------------------------------------------------------ */

void pto_sumtable(unsigned int sites, unsigned int cats,
                  const double *clvp, const double *clvc,
                  const double evecs[16], const double inv_evecs[16],
                  const double freqs[4], double *sumtable)
{
  /* LIBPLL/core_derivatives_avx.c:25-207.
     tt[j][k] = inv_evecs[k][j] * pi_k                   (:86-95)
     left_j   = dot4(tt[j], clvp), right_j = dot4(evecs[j], clvc)
     sum_j    = left_j * right_j          (scalers are not applied here) */
  double tt[16];
  for (int j = 0; j < NS; ++j)
    for (int k = 0; k < NS; ++k) tt[j * NS + k] = inv_evecs[k * NS + j] * freqs[k];
  for (size_t q = 0; q < (size_t)sites * cats; ++q)
  {
    const double *p = clvp + q * NS, *c = clvc + q * NS;
    double *s = sumtable + q * NS;
    for (int j = 0; j < NS; ++j)
      s[j] = dot4(tt + j * NS, p) * dot4(evecs + j * NS, c);
  }
}

/* ------------------------------------------------------------------ */
/* first and second derivative of -lnL at branch length t              */
/* ------------------------------------------------------------------ */

void pto_derivatives(unsigned int sites, unsigned int cats,
                     const double *sumtable, const double evals[4],
                     const double *rates, double t, double *d1, double *d2)
{
  pto_derivatives_w(sites, cats, sumtable, evals, rates, NULL, t, d1, d2);
}

void pto_derivatives_w(unsigned int sites, unsigned int cats,
                       const double *sumtable, const double evals[4],
                       const double *rates, const unsigned int *weights, double t,
                       double *d1, double *d2)
{
  /* diag table: LIBPLL/core_derivatives.c:560-575 (prop_invar = 0, so
     ki = rates[c] / (1.0 - 0.0)); site loop: LIBPLL/core_derivatives_avx.c
     :1003-1166 -- equal rate weights are dropped (:921-928,1093-1097), four
     consecutive sites share one vector step whose lanes accumulate
     independently (:1111-1142), left-over sites go through the scalar tail in
     reverse order with a true division (:1148-1156), lanes are summed last
     (:1160-1166). */
  double *diag = (double *)malloc(sizeof(double) * cats * NS * 3);
  for (unsigned int c = 0; c < cats; ++c)
  {
    const double ki = rates[c] / (1.0 - 0.0);
    for (int j = 0; j < NS; ++j)
    {
      double *dp = diag + (c * NS + j) * 3;
      dp[0] = exp(evals[j] * ki * t);
      dp[1] = evals[j] * ki * dp[0];
      dp[2] = evals[j] * ki * evals[j] * ki * dp[0];
    }
  }
  double lane1[4] = {0, 0, 0, 0}, lane2[4] = {0, 0, 0, 0};
  double blk[4][3];
  const unsigned int full = sites / 4 * 4;
  for (unsigned int n = 0; n < sites; ++n)
  {
    double lk[3] = {0, 0, 0};
    for (unsigned int c = 0; c < cats; ++c)
    {
      const double *s = sumtable + ((size_t)n * cats + c) * NS;
      double cat[3] = {0, 0, 0};
      for (int j = 0; j < NS; ++j)
        for (int q = 0; q < 3; ++q) cat[q] = cat[q] + s[j] * diag[(c * NS + j) * 3 + q];
      for (int q = 0; q < 3; ++q) lk[q] = lk[q] + cat[q];
    }
    for (int q = 0; q < 3; ++q) blk[n & 3][q] = lk[q];
    if (n < full && (n & 3) == 3)
    {
      /* four weights all 1: no products (:1123-1130), else every lane is multiplied by its
         weight (:1131-1139) */
      const int unit = !weights ||
                       (weights[n - 3] | weights[n - 2] | weights[n - 1] | weights[n]) == 1;
      for (int v = 0; v < 4; ++v)
      {
        const double recip = 1.0 / blk[v][0];
        const double a = blk[v][1] * recip;
        const double b = a * a - blk[v][2] * recip;
        const double w = unit ? 1.0 : (double)weights[n - 3 + v];
        lane1[v] = lane1[v] - (unit ? a : a * w);
        lane2[v] = lane2[v] + (unit ? b : b * w);
      }
    }
  }
  double df = 0., ddf = 0.;
  for (unsigned int n = sites; n > full; --n)
  {
    const double *lk = blk[(n - 1) & 3];
    const double a = (-lk[1] / lk[0]);
    const double b = (a * a - (lk[2] / lk[0]));
    df += (weights ? weights[n - 1] : 1u) * a;
    ddf += (weights ? weights[n - 1] : 1u) * b;
  }
  df += lane1[0] + lane1[1] + lane1[2] + lane1[3];
  ddf += lane2[0] + lane2[1] + lane2[2] + lane2[3];
  *d1 = df;
  *d2 = ddf;
  free(diag);
}

/* ------------------------------------------------------------------ */
/* pt's Newton-Raphson on one branch                                   */
/* ------------------------------------------------------------------ */

double pto_optimize_branch(unsigned int sites, unsigned int cats,
                           const double *sumtable, const double evals[4],
                           const double *rates, double len0,
                           unsigned int max_iter, double eps, double map_lambda,
                           int *iters, int *status)
{
  return pto_optimize_branch_w(sites, cats, sumtable, evals, rates, NULL, len0, max_iter, eps,
                               map_lambda, iters, status);
}

double pto_optimize_branch_w(unsigned int sites, unsigned int cats,
                             const double *sumtable, const double evals[4],
                             const double *rates, const unsigned int *weights, double len0,
                             unsigned int max_iter, double eps, double map_lambda,
                             int *iters, int *status)
{
  /* PTPLL/pll_partition.cpp:419-481 */
  double len = len0;
  int maybe_decreasing = 0, it = 0;
  *status = 0;
  for (unsigned int i = 0; i < max_iter; ++i)
  {
    double d1, d2;
    pto_derivatives_w(sites, cats, sumtable, evals, rates, weights, len, &d1, &d2);
    ++it;
    if (map_lambda != 0.0) d1 += map_lambda;
    if (!isfinite(d1)) { *status = 1; break; }
    if (!isfinite(d2)) { *status = 2; break; }
    if (fabs(d1) < eps) break;
    if (d2 < 0.0) len += d1 / d2;
    else len -= d1 / d2;
    if (len < 0.0)
    {
      len = eps;
      if (maybe_decreasing) break;
      maybe_decreasing = 1;
    }
  }
  if (iters) *iters = it;
  return len;
}

/* ------------------------------------------------------------------ */
/* site pattern compression                                            */
/* ------------------------------------------------------------------ */

static unsigned int g_cmp_count;
static int cmp_columns(const void *a, const void *b)
{
  return memcmp(*(const unsigned char *const *)a, *(const unsigned char *const *)b, g_cmp_count);
}

int pto_compress_site_patterns(char **sequence, unsigned int count, unsigned int *length,
                               unsigned int *weight)
{
  /* LIBPLL/compress.c:138-282 with map = pll_map_nt (all codes < 256, so charmap = map, :160-164).
     characters -> state codes (:173), alignment split into columns (:214-219), columns sorted
     (:222; a three-way radix quicksort on the code strings = ascending strcmp order, so any
     sort gives the same sequence of DISTINCT columns), first of every run kept and the run
     length is its weight (:225-241), codes decoded through the inverse map, which holds the
     LAST character of each code (:166-170, :244-246), terminator added (:248-249).
     returns 0, or -1 on a character outside pll_map_nt (the reference would encode it as 0
     and truncate the column string there; pt rejects such input before, PTPLL/pll_partition.cpp:184-187) */
  const unsigned int n = *length;
  unsigned char inv[256];
  memset(inv, 0, sizeof inv);
  for (int c = 1; c < 256; ++c)
    if (pto_nt_mask(c)) inv[pto_nt_mask(c)] = (unsigned char)c;
  if (!n) return 0;
  unsigned char *mem = (unsigned char *)malloc((size_t)n * count);
  unsigned char **column = (unsigned char **)malloc(sizeof(unsigned char *) * n);
  for (unsigned int i = 0; i < n; ++i) column[i] = mem + (size_t)i * count;
  for (unsigned int j = 0; j < count; ++j)
    for (unsigned int i = 0; i < n; ++i)
    {
      const unsigned int k = pto_nt_mask((unsigned char)sequence[j][i]);
      if (!k) { free(mem); free(column); return -1; }
      column[i][j] = (unsigned char)k;
    }
  g_cmp_count = count;
  qsort(column, n, sizeof column[0], cmp_columns);
  unsigned int ref = 0;
  weight[0] = 1;
  for (unsigned int i = 1; i < n; ++i)
  {
    if (memcmp(column[i], column[i - 1], count))
    {
      column[ref + 1] = column[i];
      ++ref;
      weight[ref] = 1;
    }
    else
      weight[ref]++;
  }
  const unsigned int m = ref + 1;
  for (unsigned int i = 0; i < m; ++i)
    for (unsigned int j = 0; j < count; ++j) sequence[j][i] = (char)inv[column[i][j]];
  for (unsigned int j = 0; j < count; ++j) sequence[j][m] = 0;
  *length = m;
  free(mem);
  free(column);
  return 0;
}
