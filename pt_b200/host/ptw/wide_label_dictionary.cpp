// pt_b200/host/ptw/wide_label_dictionary.cpp -- see the header.
#include "wide_label_dictionary.hpp"

#include <limits>
#include <stdexcept>

namespace pt {

LabelDictionary::LabelDictionary() {}

LabelDictionary::LabelDictionary(const std::vector<std::string>& labels)
{
  if (labels.empty()) throw std::invalid_argument("labels is empty");
  const size_t capacity = static_cast<size_t>(std::numeric_limits<index_type>::max()) + 1;
  if (labels.size() > capacity) throw std::length_error("index type is not large enough");
  labels_.reserve(labels.size());
  index_of_.reserve(2 * labels.size());
  for (const std::string& label : labels) {
    const index_type next = static_cast<index_type>(labels_.size());
    if (!index_of_.emplace(label, next).second) throw std::invalid_argument("duplicate label");
    labels_.push_back(label);
  }
}

std::string LabelDictionary::GetLabel(index_type index) const
{
  if (index >= labels_.size()) throw std::invalid_argument("index not found in dictionary");
  return labels_[index];
}

LabelDictionary::index_type LabelDictionary::GetIndex(const std::string& label) const
{
  const auto it = index_of_.find(label);
  if (it == index_of_.end()) throw std::invalid_argument("label not found in dictionary");
  return it->second;
}

} // namespace pt
