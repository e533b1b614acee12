// pt_b200/csrc/gtr_model.cpp -- host-side model set-up of the engine.
//
// The time-reversible 4x4 rate matrix is diagonalised once per SetModel on the
// host (it is 4x4; SURVEY.md section 2a keeps pll_update_eigen on the host) and
// handed to the kernels as constants.  To keep P matrices comparable with the
// reference down to the last bits the same classical algorithm is used
// (Householder tridiagonalisation + implicit-shift QL, EISPACK tred2/tql2 as
// published in Numerical Recipes), on the symmetrised matrix
// sqrt(pi) Q sqrt(pi)^-1 in the same operation order as
// /root/reference/lib/libptpll/lib/pll-modules/libs/libpll/src/models.c:24-331.
#include "gtr_model.hpp"

#include <array>
#include <cmath>

namespace ptgpu {

namespace {

using Mat4 = std::array<std::array<double, 4>, 4>;
using Vec4 = std::array<double, 4>;

constexpr int N = 4;

// reduce the symmetric matrix m to tridiagonal form; on return m holds the
// accumulated orthogonal transform, diag/off the tridiagonal (models.c:100-178)
void tridiagonalise(Mat4 &m, Vec4 &diag, Vec4 &off)
{
  for (int col = N - 1; col > 0; --col)
  {
    const int rows = col;
    double h = 0.0;
    if (rows > 1)
    {
      double norm1 = 0.0;
      for (int k = 0; k < rows; ++k) norm1 += std::fabs(m[k][col]);
      if (norm1 == 0.0)
        off[col] = m[rows - 1][col];
      else
      {
        for (int k = 0; k < rows; ++k)
        {
          m[k][col] /= norm1;
          h += m[k][col] * m[k][col];
        }
        double f = m[rows - 1][col];
        double g = (f > 0) ? -std::sqrt(h) : std::sqrt(h);
        off[col] = norm1 * g;
        h -= f * g;
        m[rows - 1][col] = f - g;
        f = 0.0;
        for (int j = 0; j < rows; ++j)
        {
          m[col][j] = m[j][col] / h;
          g = 0.0;
          for (int k = 0; k <= j; ++k) g += m[k][j] * m[k][col];
          for (int k = j + 1; k < rows; ++k) g += m[j][k] * m[k][col];
          off[j] = g / h;
          f += off[j] * m[j][col];
        }
        const double hh = f / (h + h);
        for (int j = 0; j < rows; ++j)
        {
          f = m[j][col];
          g = off[j] - hh * f;
          off[j] = g;
          for (int k = 0; k <= j; ++k) m[k][j] -= (f * off[k] + g * m[k][col]);
        }
      }
    }
    else
      off[col] = m[rows - 1][col];
    diag[col] = h;
  }
  diag[0] = 0.0;
  off[0] = 0.0;
  for (int col = 0; col < N; ++col)
  {
    if (diag[col] != 0.0)
      for (int j = 0; j < col; ++j)
      {
        double g = 0.0;
        for (int k = 0; k < col; ++k) g += m[k][col] * m[j][k];
        for (int k = 0; k < col; ++k) m[j][k] -= g * m[col][k];
      }
    diag[col] = m[col][col];
    m[col][col] = 1.0;
    for (int j = 0; j < col; ++j) m[col][j] = m[j][col] = 0.0;
  }
}

// QL with implicit shifts; rows of m end up as eigenvectors (models.c:24-97)
void ql_rows(Vec4 &diag, Vec4 &off, Mat4 &m)
{
  for (int i = 1; i < N; ++i) off[i - 1] = off[i];
  off[N - 1] = 0.0;
  for (int lo = 0; lo < N; ++lo)
  {
    for (;;)
    {
      int hi = lo;
      for (; hi < N - 1; ++hi)
      {
        const double dd = std::fabs(diag[hi]) + std::fabs(diag[hi + 1]);
        if (std::fabs(off[hi]) + dd == dd) break;
      }
      if (hi == lo) break;
      double g = (diag[lo + 1] - diag[lo]) / (2.0 * off[lo]);
      double r = std::sqrt((g * g) + 1.0);
      g = diag[hi] - diag[lo] + off[lo] / (g + ((g < 0) ? -std::fabs(r) : std::fabs(r)));
      double sn = 1.0, cs = 1.0, p = 0.0;
      for (int i = hi - 1; i >= lo; --i)
      {
        double f = sn * off[i];
        const double b = cs * off[i];
        if (std::fabs(f) >= std::fabs(g))
        {
          cs = g / f;
          r = std::sqrt((cs * cs) + 1.0);
          off[i + 1] = f * r;
          sn = 1.0 / r;
          cs *= sn;
        }
        else
        {
          sn = f / g;
          r = std::sqrt((sn * sn) + 1.0);
          off[i + 1] = g * r;
          cs = 1.0 / r;
          sn *= cs;
        }
        g = diag[i + 1] - p;
        r = (diag[i] - g) * sn + 2.0 * cs * b;
        p = sn * r;
        diag[i + 1] = g + p;
        g = cs * r - b;
        for (int k = 0; k < N; ++k)
        {
          f = m[i + 1][k];
          m[i + 1][k] = sn * m[i][k] + cs * f;
          m[i][k] = cs * m[i][k] - sn * f;
        }
      }
      diag[lo] = diag[lo] - p;
      off[lo] = g;
      off[hi] = 0.0;
    }
  }
}

} // namespace

void gtr_eigensystem(const double freqs[4], const double subst[6], double evecs[16],
                     double inv_evecs[16], double evals[4])
{
  // models.c:182-249
  double rate[6];
  for (int i = 0; i < 6; ++i) rate[i] = subst[i];
  if (rate[5] > 0.0)
    for (int i = 0; i < 6; ++i) rate[i] /= rate[5];

  Mat4 q{};
  int k = 0;
  for (int i = 0; i < N; ++i)
    for (int j = i + 1; j < N; ++j)
    {
      const double factor = rate[k++];
      q[i][j] = q[j][i] = factor * std::sqrt(freqs[i] * freqs[j]);
      q[i][i] -= factor * freqs[j];
      q[j][j] -= factor * freqs[i];
    }
  double mean = 0;
  for (int i = 0; i < N; ++i) mean += freqs[i] * (-q[i][i]);
  for (auto &row : q)
    for (double &x : row) x /= mean;

  Vec4 diag{}, off{};
  tridiagonalise(q, diag, off);
  ql_rows(diag, off, q);

  // models.c:290-318: V = rows * sqrt(pi_j), V^-1 = rows^T / sqrt(pi_i)
  for (int i = 0; i < N; ++i)
  {
    evals[i] = diag[i];
    for (int j = 0; j < N; ++j)
    {
      inv_evecs[i * N + j] = q[j][i] / std::sqrt(freqs[i]);
      evecs[i * N + j] = q[i][j] * std::sqrt(freqs[j]);
    }
  }
}

unsigned int nucleotide_mask(unsigned char ch)
{
  // IUPAC nucleotide codes as 4-bit sets, bit0=A bit1=C bit2=G bit3=T
  // (pll_map_nt, libpll maps.c:46-64); 0 = not a legal state
  if (ch >= 'a' && ch <= 'z') ch = (unsigned char)(ch - 'a' + 'A');
  static const char *codes = "ACMGRSVTWYHKDBN"; // index+1 == mask
  for (unsigned int i = 0; codes[i]; ++i)
    if (codes[i] == ch) return i + 1;
  switch (ch)
  {
    case 'U': return 8;
    case 'O': case 'X': case '-': case '?': return 15;
    default: return 0;
  }
}

} // namespace ptgpu

// ---------------------------------------------------------------------------
// Discrete-Gamma category rates (mean of each equal-probability category),
// Yang (1994) J Mol Evol 39:306-314, built from the published routines it
// cites: Lanczos/Stirling ln Gamma (Pike & Hill 1966, ACM 291), incomplete
// gamma ratio (Bhattacharjee 1970, AS 32), normal quantile (Odeh & Evans 1974,
// AS 70) and chi-square quantile (Best & Roberts 1975, AS 91).  The operation
// order is that of the reference's host routine
// (/root/reference/lib/libptpll/lib/pll-modules/libs/libpll/src/gamma.c:24-292)
// because the rates feed every P matrix: they have to be the same doubles.
// ---------------------------------------------------------------------------
namespace ptgpu {
namespace {

double log_gamma(double a)
{
  double x = a, lead = 0.0;
  if (x < 7.0)
  {
    double prod = 1.0, z = a - 1.0;
    while ((z = z + 1.0) < 7.0) prod *= z;
    x = z;
    lead = -std::log(prod);
  }
  const double z = 1 / (x * x);
  return lead + (x - 0.5) * std::log(x) - x + .918938533204673 +
         (((-.000595238095238 * z + .000793650793651) * z - .002777777777778) * z +
          .083333333333333) / x;
}

// P(alpha, x) = regularised lower incomplete gamma; -1 on a domain error
double gamma_ratio(double x, double alpha, double lg_alpha)
{
  constexpr double kAccurate = 1e-8, kOverflow = 1e30;
  if (x == 0) return 0;
  if (x < 0 || alpha <= 0) return -1;
  const double front = std::exp(alpha * std::log(x) - x - lg_alpha);
  if (!(x > 1 && x >= alpha))
  {
    double sum = 1, term = 1, k = alpha;
    do
    {
      k++;
      term *= x / k;
      sum += term;
    } while (term > kAccurate);
    sum *= front / alpha;
    return sum;
  }
  // continued fraction with renormalisation of the convergents
  double a = 1 - alpha, b = a + x + 1, n = 0;
  double conv[6] = {1, x, x + 1, x * b, 0, 0};
  double ratio = conv[2] / conv[3];
  for (;;)
  {
    a++;
    b += 2;
    n++;
    const double an = a * n;
    conv[4] = b * conv[2] - an * conv[0];
    conv[5] = b * conv[3] - an * conv[1];
    if (conv[5] != 0)
    {
      const double next = conv[4] / conv[5];
      const double gap = std::fabs(ratio - next);
      if (!(gap > kAccurate) && gap <= kAccurate * next) return 1 - front * ratio;
      ratio = next;
    }
    for (int i = 0; i < 4; ++i) conv[i] = conv[i + 2];
    if (!(std::fabs(conv[4]) < kOverflow))
      for (int i = 0; i < 4; ++i) conv[i] /= kOverflow;
  }
}

double normal_quantile(double p)
{
  const double q = p < 0.5 ? p : 1 - p;
  if (q < 1e-20) return -9999;
  const double y = std::sqrt(std::log(1 / (q * q)));
  const double num = (((y * -.453642210148e-4 + -.0204231210245) * y + -.342242088547) * y + -1) * y +
                     -.322232431088;
  const double den = (((y * .0038560700634 + .103537752850) * y + .531103462366) * y + .588581570495) * y +
                     .0993484626060;
  const double z = y + num / den;
  return p < 0.5 ? -z : z;
}

double chi2_quantile(double p, double v)
{
  constexpr double kEps = .5e-6, kLn2 = .6931471805;
  if (p < .000002 || p > .999998 || v <= 0) return -1;
  const double g = log_gamma(v / 2);
  const double half = v / 2, c = half - 1;
  double ch;
  if (!(v >= -1.24 * std::log(p)))
  {
    ch = std::pow(p * half * std::exp(g + half * kLn2), 1 / half);
    if (ch - kEps < 0) return ch;
  }
  else if (v > .32)
  {
    const double x = normal_quantile(p);
    const double p1 = 0.222222 / v;
    ch = v * std::pow(x * std::sqrt(p1) + 1 - p1, 3.0);
    if (ch > 2.2 * v + 6) ch = -2 * (std::log(1 - p) - c * std::log(.5 * ch) + g);
  }
  else
  {
    ch = 0.4;
    const double a = std::log(1 - p);
    double prev;
    do
    {
      prev = ch;
      const double p1 = 1 + ch * (4.67 + ch);
      const double p2 = ch * (6.73 + ch * (6.66 + ch));
      const double t = -0.5 + (4.67 + 2 * ch) / p1 - (6.73 + ch * (13.32 + 3 * ch)) / p2;
      ch -= (1 - std::exp(a + g + .5 * ch + c * kLn2) * p2 / p1) / t;
    } while (!(std::fabs(prev / ch - 1) - .01 <= 0));
  }
  double prev;
  do
  {
    prev = ch;
    const double p1 = .5 * ch;
    double t = gamma_ratio(p1, half, g);
    if (t < 0.0) return -1;
    const double p2 = p - t;
    t = p2 * std::exp(half * kLn2 + g + p1 - c * std::log(ch));
    const double b = t / ch;
    const double a = 0.5 * t - b * c;
    const double s1 = (210 + a * (140 + a * (105 + a * (84 + a * (70 + 60 * a))))) / 420;
    const double s2 = (420 + a * (735 + a * (966 + a * (1141 + 1278 * a)))) / 2520;
    const double s3 = (210 + a * (462 + a * (707 + 932 * a))) / 2520;
    const double s4 = (252 + a * (672 + 1182 * a) + c * (294 + a * (889 + 1740 * a))) / 5040;
    const double s5 = (84 + 264 * a + c * (175 + 606 * a)) / 2520;
    const double s6 = (120 + c * (346 + 127 * c)) / 5040;
    ch += t * (1 + 0.5 * t * s1 - b * c * (s1 - b * (s2 - b * (s3 - b * (s4 - b * (s5 - b * s6))))));
  } while (std::fabs(prev / ch - 1) > kEps);
  return ch;
}

} // namespace

bool gamma_mean_rates(double alpha, unsigned int cats, double *rates)
{
  if (alpha < 0.02 || cats < 1 || !rates) return false;
  if (cats == 1) { rates[0] = 1.0; return true; }
  const double scale = alpha / alpha * cats;
  const double lg1 = log_gamma(alpha + 1);
  std::array<double, 64> cut{};
  if (cats > cut.size()) return false;
  for (unsigned int i = 0; i + 1 < cats; ++i)
    cut[i] = chi2_quantile((i + 1.0) / cats, 2.0 * alpha) / (2.0 * alpha);
  for (unsigned int i = 0; i + 1 < cats; ++i) cut[i] = gamma_ratio(cut[i] * alpha, alpha + 1, lg1);
  rates[0] = cut[0] * scale;
  rates[cats - 1] = (1 - cut[cats - 2]) * scale;
  for (unsigned int i = 1; i + 1 < cats; ++i) rates[i] = (cut[i] - cut[i - 1]) * scale;
  return true;
}

} // namespace ptgpu
