// pt_b200/host/ptw/compressed_tree.cpp -- pt::CompressedTree (the topology key of the authority's
// three tree tables) without the whole-tree clone and the string comparisons.
//
// An implementation of the class declared in pt's own src/compressed_tree.hpp, linked into
// ptw_batched IN PLACE of the reference's src/compressed_tree.cpp; the class layout, the key's
// encoding and therefore equality, hashing and the printed Newick strings are the reference's:
//
//   reference (src/compressed_tree.cpp:79-146, src/ordered_tree.cpp:118-187), per key:
//     pll_utree_clone of the whole tree; FindRootNode three times over it, comparing and copying
//     std::string labels at every inner node and re-linking the clone's records into order; a
//     std::map<std::string, index> lookup per tip; pll_utree_destroy.
//   here: one pass over the tips (label -> dictionary index through a per-thread table that is
//     verified with one strcmp per tip, alphabetical RANKS precomputed once per dictionary), one
//     post-order pass of integer minima over the records, one pre-order pass that emits the bits.
//     The caller's tree is only read.
//
// The key every Authority::ProposeMove / RequestMove / ReportTestScore / ReportVisitScore builds
// (src/authority.cpp:124-214) is what the walk's host time consists of once the likelihood work is
// batched and computed ahead of time.
//
// Canonical form, as the reference defines it: reroot at the parent of the alphabetically smallest
// tip; that tip is the first subtree; at every other node the child subtree holding the smaller
// minimum label comes first; pre-order, one bit per node (1 = inner, 0 = tip), tips carry their
// dictionary index.  (The index type is the reference's LabelDictionary::index_type, uint8_t:
// the 256-taxon limit is part of pt's header and stays.)
#include "compressed_tree.hpp"

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <functional>
#include <numeric>
#include <sstream>
#include <stdexcept>

#include <libpll/pll.h>

namespace pt {

namespace {

// alphabetical rank of every dictionary index (std::string order == the reference's
// label.compare()), rebuilt by BuildDictionary; `generation` invalidates the per-thread tables
std::vector<unsigned int> g_rank_of_index;
std::vector<std::string> g_label_of_index;
std::atomic<unsigned long long> g_generation{1};

struct TipEntry {
  unsigned int index = 0;
  bool valid = false;
};
struct TipTable {
  unsigned long long generation = 0;
  std::vector<TipEntry> by_node_index;
};

} // namespace

LabelDictionary CompressedTree::label_dictionary_ = LabelDictionary();

void CompressedTree::BuildDictionary(const std::vector<std::string>& labels)
{
  label_dictionary_ = LabelDictionary(labels);
  std::vector<unsigned int> order(labels.size());
  std::iota(order.begin(), order.end(), 0u);
  std::sort(order.begin(), order.end(), [&labels](unsigned int a, unsigned int b) { return labels[a] < labels[b]; });
  g_rank_of_index.assign(labels.size(), 0u);
  for (unsigned int r = 0; r < order.size(); ++r) g_rank_of_index[order[r]] = r;
  g_label_of_index = labels;
  g_generation.fetch_add(1);
}

CompressedTree::CompressedTree(const pll_utree_t* tree)
{
  Encode(tree);
}

void CompressedTree::Encode(const pll_utree_t* tree)
{
  if (!tree) throw std::invalid_argument("tree is null");
  const unsigned int tips = tree->tip_count, inner = tree->inner_count;
  const unsigned int records = tips + 3 * inner;

  thread_local TipTable tls_table;
  TipTable& table = tls_table;   // (one TLS lookup, not one per use)
  const unsigned long long generation = g_generation.load(std::memory_order_relaxed);
  if (table.generation != generation) {
    table.by_node_index.clear();
    table.generation = generation;
  }
  if (table.by_node_index.size() < records) table.by_node_index.resize(records);

  // rank (alphabetical position of the smallest label below) per record, dictionary index per tip
  thread_local std::vector<unsigned int> tls_rank, tls_dict_index;
  tls_rank.assign(records, 0u);
  tls_dict_index.assign(records, 0u);
  unsigned int* const rank = tls_rank.data();
  unsigned int* const dict_index = tls_dict_index.data();
  const pll_unode_t* smallest = nullptr;
  for (unsigned int i = 0; i < tips; ++i) {
    const pll_unode_t* tip = tree->nodes[i];
    if (tip->next || !tip->label || tip->node_index >= records)
      throw std::invalid_argument("tree has an unlabelled or misplaced tip");
    TipEntry& entry = table.by_node_index[tip->node_index];
    if (!entry.valid || std::strcmp(tip->label, g_label_of_index[entry.index].c_str()) != 0) {
      entry.index = label_dictionary_.GetIndex(tip->label);   // throws for an unknown label
      entry.valid = true;
    }
    dict_index[tip->node_index] = entry.index;
    rank[tip->node_index] = g_rank_of_index[entry.index];
    if (!smallest || rank[tip->node_index] < rank[smallest->node_index]) smallest = tip;
  }
  if (!smallest) throw std::invalid_argument("root is null");
  const pll_unode_t* root = smallest->back;
  if (!root || !root->next) throw std::invalid_argument("Encode() requires an internal node");

  // The authority builds the key of the same tree two or three times in a row (ProposeMove,
  // ReportTestScore, RequestMove on one move: src/authority.cpp:124-183, src/wanderer.cpp:267-278).
  // What the key depends on -- which record is linked to which, which label sits on which tip --
  // is summed into two independent 64-bit hashes; the last few keys are kept per thread under them.
  struct Memo {
    uint64_t h1 = 0, h2 = 0;
    unsigned long long generation = 0;
    std::vector<bool> bits;
    std::vector<index_type> labels;
  };
  constexpr int kMemo = 4;
  thread_local Memo tls_memo[kMemo];
  thread_local int tls_memo_next = 0;
  Memo* const memo = tls_memo;
  int& memo_next = tls_memo_next;
  uint64_t h1 = records, h2 = tips;
  {
    auto mix = [&h1, &h2](uint64_t a, uint64_t b) {
      uint64_t x = (a + 1) * 0x9e3779b97f4a7c15ull ^ (b + 1) * 0xc2b2ae3d27d4eb4full;
      x ^= x >> 29; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 32;
      h1 += x;
      uint64_t y = (a + 1) * 0xd6e8feb86659fd93ull + (b + 1) * 0x165667b19e3779f9ull;
      y ^= y >> 31; y *= 0x94d049bb133111ebull; y ^= y >> 30;
      h2 += y;
    };
    for (unsigned int i = 0; i < tips; ++i) {
      const pll_unode_t* tip = tree->nodes[i];
      mix(tip->node_index, tip->back->node_index);
      mix(records + tip->node_index, dict_index[tip->node_index]);
    }
    for (unsigned int i = tips; i < tips + inner; ++i) {
      const pll_unode_t* r = tree->nodes[i];
      mix(r->node_index, r->back->node_index);
      mix(r->next->node_index, r->next->back->node_index);
      mix(r->next->next->node_index, r->next->next->back->node_index);
      // the ring itself (which three records form a node)
      const unsigned int lo = std::min(r->node_index, std::min(r->next->node_index, r->next->next->node_index));
      mix(2ull * records + r->node_index, lo);
      mix(2ull * records + r->next->node_index, lo);
      mix(2ull * records + r->next->next->node_index, lo);
    }
    for (int k = 0; k < kMemo; ++k)
      if (memo[k].generation == generation && memo[k].h1 == h1 && memo[k].h2 == h2) {
        bits_ = memo[k].bits;
        label_indices_ = memo[k].labels;
        return;
      }
  }

  // post-order: minimum rank below every record that hangs under the new root
  struct Frame { const pll_unode_t* r; bool expanded; };
  thread_local std::vector<Frame> tls_stack;
  std::vector<Frame>& stack = tls_stack;
  stack.clear();
  stack.push_back(Frame{root->next->back, false});
  stack.push_back(Frame{root->next->next->back, false});
  while (!stack.empty()) {
    const Frame f = stack.back();
    stack.pop_back();
    if (!f.r->next) continue;   // tips carry their own rank
    const pll_unode_t* a = f.r->next->back;
    const pll_unode_t* b = f.r->next->next->back;
    if (!f.expanded) {
      stack.push_back(Frame{f.r, true});
      stack.push_back(Frame{a, false});
      stack.push_back(Frame{b, false});
    } else {
      rank[f.r->node_index] = std::min(rank[a->node_index], rank[b->node_index]);
    }
  }

  // pre-order emission, smaller minimum first
  bits_.clear();
  label_indices_.clear();
  bits_.reserve(tips + inner + 1);
  label_indices_.reserve(tips);
  bits_.push_back(true);
  bits_.push_back(false);
  label_indices_.push_back(static_cast<index_type>(dict_index[smallest->node_index]));
  thread_local std::vector<const pll_unode_t*> tls_todo;
  std::vector<const pll_unode_t*>& todo = tls_todo;
  todo.clear();
  {
    const pll_unode_t* a = root->next->back;
    const pll_unode_t* b = root->next->next->back;
    if (rank[b->node_index] < rank[a->node_index]) std::swap(a, b);
    todo.push_back(b);
    todo.push_back(a);
  }
  while (!todo.empty()) {
    const pll_unode_t* r = todo.back();
    todo.pop_back();
    if (!r->next) {
      bits_.push_back(false);
      label_indices_.push_back(static_cast<index_type>(dict_index[r->node_index]));
      continue;
    }
    bits_.push_back(true);
    const pll_unode_t* a = r->next->back;
    const pll_unode_t* b = r->next->next->back;
    if (rank[b->node_index] < rank[a->node_index]) std::swap(a, b);
    todo.push_back(b);
    todo.push_back(a);
  }
  Memo& slot = memo[memo_next];
  memo_next = (memo_next + 1) % kMemo;
  slot.h1 = h1;
  slot.h2 = h2;
  slot.generation = generation;
  slot.bits = bits_;
  slot.labels = label_indices_;
}

// the Newick string the reference prints for a decoded key (src/ordered_tree.cpp:16-84:
// "(a,b,c);" at the root, "(a,b)" below, labels at the tips), straight from the bits
std::string CompressedTree::Decode() const
{
  if (bits_.empty()) throw std::invalid_argument("bits is empty");
  if (!bits_[0]) throw std::invalid_argument("first bit is not set");
  std::string out;
  size_t bit = 1, label = 0;
  std::function<void()> subtree = [&]() {
    if (bit >= bits_.size()) throw std::runtime_error("bits is empty");
    if (bits_[bit++]) {
      out += '(';
      subtree();
      out += ',';
      subtree();
      out += ')';
    } else {
      if (label >= label_indices_.size()) throw std::runtime_error("label indices are empty");
      const std::string name = label_dictionary_.GetLabel(label_indices_[label++]);
      if (name.empty()) throw std::invalid_argument("empty tip node label");
      out += name;
    }
  };
  out += '(';
  subtree();
  out += ',';
  subtree();
  out += ',';
  subtree();
  out += ");";
  return out;
}

std::string CompressedTree::ToDebugString() const
{
  std::stringstream ss;
  ss << "[";
  for (auto bit : bits_) ss << (bit ? "1" : "0");
  ss << "]\n";
  std::string delim = "";
  ss << "[";
  for (auto label_index : label_indices_) {
    ss << delim << static_cast<unsigned int>(label_index);
    delim = ", ";
  }
  ss << "]\n";
  return ss.str();
}

// same combination as the reference's (src/compressed_tree.cpp:246-257), so that the tables
// iterate -- and the walk's output files list their trees -- in the reference's order
size_t CompressedTree::Hash() const
{
  size_t seed = std::hash<std::vector<bool>>()(bits_);
  for (auto label_index : label_indices_)
    seed ^= std::hash<index_type>()(label_index) + 0x9e3779b9 + (seed << 6) + (seed >> 2);
  return seed;
}

} // namespace pt
