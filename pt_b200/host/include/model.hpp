// pt_b200/host/include/model.hpp -- substitution-model record of the host layer.
//
// Mirrors the interface of the reference's lib/libptpll/src/model.hpp:10-23
// (struct Model, ParseRaxmlInfo, comparison / stream operators) so callers of
// pt::pll::Partition compile unchanged.
#ifndef PT_B200_HOST_MODEL_HPP_
#define PT_B200_HOST_MODEL_HPP_

#include <cstddef>
#include <iosfwd>
#include <string>
#include <vector>

namespace pt { namespace pll {

struct Model {
  std::string model_name;               // "GTR", "JC", "HKY", ... (pll-modules DNA model names)
  std::vector<double> frequencies;      // pi_A pi_C pi_G pi_T
  std::vector<double> subst_params;     // ac ag at cg ct gt
  std::vector<double> category_rates;   // discrete-Gamma mean rates
};

// Reads "frequencies:", "ac ag at cg ct gt:", "alpha[0]:" and "Substitution
// Matrix:" out of a RAxML_info file; category rates are the Gamma(alpha) mean
// rates for `rate_categories` categories (RAxML's default is 4).
Model ParseRaxmlInfo(const std::string& path, size_t rate_categories = 4);

std::ostream& operator<<(std::ostream& os, const Model& rhs);
bool operator==(const Model& lhs, const Model& rhs);
bool operator!=(const Model& lhs, const Model& rhs);

} } // namespace pt::pll

#endif // PT_B200_HOST_MODEL_HPP_
