// pt_b200/host/ptw/wide_label_dictionary.hpp -- pt::LabelDictionary with a 16-bit label index
// ("a label index wider than uint8_t").
//
// pt's own class (src/label_dictionary.hpp:10-26) fixes `index_type = uint8_t`, and
// pt::CompressedTree -- the key of the authority's three tree tables -- stores one index per
// tip (src/compressed_tree.hpp:17-22): a tree of more than 256 taxa cannot be walked
// (src/label_dictionary.cpp:19-22 throws std::length_error).  The type is part of the class
// layout, so it cannot be widened from outside a translation unit; it CAN be widened for a
// whole build: this header is passed with `-include` to every pt source that goes into
// ptw_batched (pt_b200/host/Makefile, the `ptwide` objects).  It claims the include guard of
// src/label_dictionary.hpp, so pt's own header contributes nothing when compressed_tree.hpp
// includes it, and declares the same public interface with `index_type = uint16_t` (65 536
// taxa).  Implementation: wide_label_dictionary.cpp (linked in place of src/label_dictionary.cpp).
//
// Nothing else changes: for dictionaries of up to 256 labels the keys, their hashes (a
// std::hash of an unsigned integer is its value, whatever the width) and the printed Newick
// strings are identical to the reference's -- the DS1 and stand-in walk tables are compared
// against the reference's in tests/test_gpu_dropin.py / test_gpu_standins.py with this build.
#ifndef PT_LABEL_DICTIONARY_HPP_
#define PT_LABEL_DICTIONARY_HPP_

#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

namespace pt {

class LabelDictionary {
 public:
  using index_type = uint16_t;

  LabelDictionary();
  // labels[i] gets index i; throws std::invalid_argument for an empty list or a repeated
  // label, std::length_error beyond 65 536 labels (the reference's conditions)
  explicit LabelDictionary(const std::vector<std::string>& labels);

  std::string GetLabel(index_type index) const;        // std::invalid_argument if unknown
  index_type GetIndex(const std::string& label) const;  // std::invalid_argument if unknown

 private:
  std::vector<std::string> labels_;                         // by index
  std::unordered_map<std::string, index_type> index_of_;
};

} // namespace pt

#endif /* PT_LABEL_DICTIONARY_HPP_ */
