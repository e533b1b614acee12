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
