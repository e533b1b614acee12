// pt_b200/csrc/gtr_model.hpp -- host-side model helpers of the engine.
#pragma once

namespace ptgpu {

// Eigensystem of the GTR rate matrix (freqs pi, rates ac ag at cg ct gt):
// evecs[j*4+k], inv_evecs[i*4+k], evals[k] with P(t) = I + (V^-1 o expm1(L t)) V.
void gtr_eigensystem(const double freqs[4], const double subst[6], double evecs[16],
                     double inv_evecs[16], double evals[4]);

// IUPAC nucleotide code -> 4-bit state set (A=1 C=2 G=4 T=8), 0 if illegal.
unsigned int nucleotide_mask(unsigned char ch);

// Mean rates of `cats` equal-probability categories of Gamma(alpha, alpha)
// (Yang 1994); false if alpha < 0.02 or cats is 0 / > 64.
bool gamma_mean_rates(double alpha, unsigned int cats, double *rates);

} // namespace ptgpu
