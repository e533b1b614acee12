// pt_b200/host/src/model.cpp -- RAxML_info reader and Model operators.
// Behaviour follows the reference's lib/libptpll/src/model.cpp:27-145; the
// Gamma category rates come from the engine library's host routine
// (ptgpu_gamma_rates_mean) instead of libpll's pll_compute_gamma_cats.
#include "model.hpp"

#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include "ptgpu.h"

namespace pt { namespace pll {

namespace {

// text after `key` up to (not including) the first `stop` character
std::string field_after(const std::string& text, const std::string& key, size_t skip, char stop)
{
  const size_t at = text.find(key);
  const size_t end = text.find(stop, stop == ' ' ? at + skip : at);
  return text.substr(at + skip, end - at - skip);
}

std::vector<double> numbers(const std::string& s)
{
  std::vector<double> out;
  std::stringstream ss(s);
  std::string tok;
  while (std::getline(ss, tok, ' ')) out.push_back(std::stod(tok));
  return out;
}

} // namespace

Model ParseRaxmlInfo(const std::string& path, size_t rate_categories)
{
  std::ifstream file(path);
  if (!file) throw std::invalid_argument("Error opening file " + path);
  std::stringstream buf;
  std::string line;
  while (std::getline(file, line)) buf << line << '\n';
  const std::string text = buf.str();

  Model m;
  m.frequencies = numbers(field_after(text, "frequencies: ", 13, '\n'));
  m.subst_params = numbers(field_after(text, "ac ag at cg ct gt:", 19, '\n'));
  if (m.subst_params.size() != (m.frequencies.size() * m.frequencies.size() - 4) / 2)
    throw std::invalid_argument("Wrong number of rate values.");

  const double alpha = std::stod(field_after(text, "alpha[0]: ", 10, ' '));
  m.category_rates.resize(rate_categories);
  if (ptgpu_gamma_rates_mean(alpha, static_cast<unsigned int>(rate_categories),
                             m.category_rates.data()) != PTGPU_OK)
    throw std::invalid_argument(std::string("Cannot compute Gamma category rates: ") +
                                ptgpu_last_error());

  m.model_name = field_after(text, "Substitution Matrix: ", 21, '\n');
  // RAxML names -> pll-modules names
  if (m.model_name == "JC69") m.model_name = "JC";
  else if (m.model_name == "HKY85") m.model_name = "HKY";
  return m;
}

namespace {
void put(std::ostream& os, const char* tag, const std::vector<double>& v)
{
  os << tag;
  const char* sep = "";
  for (double x : v) { os << sep << x; sep = ", "; }
  os << "\n";
}
} // namespace

std::ostream& operator<<(std::ostream& os, const Model& rhs)
{
  put(os, "freqs:  ", rhs.frequencies);
  put(os, "params: ", rhs.subst_params);
  put(os, "rates:  ", rhs.category_rates);
  return os;
}

bool operator==(const Model& lhs, const Model& rhs)
{
  return lhs.model_name == rhs.model_name && lhs.frequencies == rhs.frequencies &&
         lhs.subst_params == rhs.subst_params && lhs.category_rates == rhs.category_rates;
}

bool operator!=(const Model& lhs, const Model& rhs) { return !(lhs == rhs); }

} } // namespace pt::pll
