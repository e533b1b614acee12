// oracle/keys/keys_dump.cpp -- TEST INFRASTRUCTURE.  Reads Newick trees (one per line, same tip
// labels) and prints, per tree, pt::CompressedTree's debug string, hash and decoded Newick.
// Built twice by oracle/Makefile.keys: against the reference's src/compressed_tree.cpp and
// against pt_b200/host/ptw/compressed_tree.cpp; tests/test_keys_cpu.py requires identical output.
// With a second argument N it also times N key constructions per tree (a third argument K: K in a row).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <libpll/pll.h>

#include "compressed_tree.hpp"

int main(int argc, char** argv)
{
  if (argc < 2) return 2;
  std::ifstream in(argv[1]);
  const int reps = argc > 2 ? std::atoi(argv[2]) : 0;
  const int per_tree = argc > 3 ? std::atoi(argv[3]) : 1;   // keys built of each tree in a row
  std::string line;
  std::vector<pll_utree_t*> trees;
  while (std::getline(in, line)) {
    if (line.empty()) continue;
    pll_utree_t* t = pll_utree_parse_newick_string(line.c_str());
    if (!t) { std::fprintf(stderr, "bad tree: %s\n", line.c_str()); return 1; }
    trees.push_back(t);
  }
  if (trees.empty()) return 1;
  std::vector<std::string> labels;
  for (unsigned int i = 0; i < trees[0]->tip_count; ++i) labels.push_back(trees[0]->nodes[i]->label);
  pt::CompressedTree::BuildDictionary(labels);
  // every tree twice in a row, then all of them once more: the authority asks for the key of one
  // tree several times running (an implementation may remember its last answers)
  for (int pass = 0; pass < 2; ++pass)
    for (pll_utree_t* t : trees)
      for (int again = 0; again < (pass == 0 ? 2 : 1); ++again) {
        pt::CompressedTree key(t);
        std::cout << key.ToDebugString() << key.Hash() << "\n" << key.Decode() << "\n";
      }
  if (reps > 0) {
    const auto t0 = std::chrono::steady_clock::now();
    size_t sink = 0;
    for (int r = 0; r < reps; ++r)
      for (pll_utree_t* t : trees)
        for (int again = 0; again < per_tree; ++again) sink += pt::CompressedTree(t).Hash();
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::fprintf(stderr, "%.3f us per key (%zu trees x %d, sink %zu)\n", 1e6 * s / (reps * trees.size() * per_tree),
                 trees.size(), reps, sink);
  }
  return 0;
}
