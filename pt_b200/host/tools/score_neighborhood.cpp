// pt_b200/host/tools/score_neighborhood.cpp -- the north-star call through the C++ surface:
//   score_neighborhood <newick> <fasta> <raxml_info> [radius] [repeats]
// builds a pt::pll::Partition exactly as pt's Guru does (src/guru.cpp:77-100), visits the tree
// (FULL TraversalUpdate + LogLikelihood) and calls Partition::ScoreNniNeighborhood -- what
// BatchedNeighborhood::EvaluateMove calls once per tree -- `repeats` times; prints one JSON line with
// the wall time per call (host clock around the C++ call, results on the host when it returns), the
// tree's lnL and the sum of the candidates' scores, so that bench.py can set it beside the same
// workload driven through the C ABI from Python.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <string>
#include <vector>

#include <libpll/pll.h>
#include <model.hpp>
#include <pll_partition.hpp>
#include <pll_util.hpp>

int main(int argc, char** argv)
{
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <newick> <fasta> <raxml_info> [radius] [repeats]\n", argv[0]);
    return 2;
  }
  try {
    const int radius = argc > 4 ? std::atoi(argv[4]) : 0;
    const int repeats = argc > 5 ? std::atoi(argv[5]) : 3;
    std::vector<pll_utree_t*> trees = pt::pll::ParseMultiNewick(argv[1]);
    if (trees.empty()) throw std::invalid_argument("no tree");
    pll_utree_t* tree = trees[0];
    std::vector<std::string> labels, sequences;
    pt::pll::ParseFasta(argv[2], tree->tip_count, labels, sequences);
    const pt::pll::Model model = pt::pll::ParseRaxmlInfo(argv[3], 4);
    pt::pll::Partition partition(tree, model, labels, sequences);
    pll_unode_t* root = tree->nodes[tree->tip_count + tree->inner_count - 1];
    partition.TraversalUpdate(root, pt::pll::TraversalType::FULL);
    const double lnl = partition.LogLikelihood(root);
    std::vector<pt::pll::Partition::NniScore> scores = partition.ScoreNniNeighborhood(tree, radius);   // warm-up
    double best = 1e300, total = 0.0;
    for (int r = 0; r < repeats; ++r) {
      const auto t0 = std::chrono::steady_clock::now();
      scores = partition.ScoreNniNeighborhood(tree, radius);
      const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      best = s < best ? s : best;
      total += s;
    }
    double sum = 0.0;
    for (const auto& m : scores) sum += m.lnl;
    std::printf("{\"candidates\": %zu, \"sites\": %zu, \"taxa\": %u, \"radius\": %d, \"repeats\": %d, "
                "\"ms_per_call_mean\": %.4f, \"ms_per_call_best\": %.4f, \"tree_lnl\": %.17g, \"lnl_sum\": %.17g, "
                "\"site_shards\": %u}\n",
                scores.size(), sequences[0].size(), tree->tip_count, radius, repeats, 1e3 * total / repeats,
                1e3 * best, lnl, sum, partition.shard_count());
    for (pll_utree_t* t : trees) pll_utree_destroy(t, pt::pll::cb_erase_data);
    return 0;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "score_neighborhood: %s\n", e.what());
    return 1;
  }
}
