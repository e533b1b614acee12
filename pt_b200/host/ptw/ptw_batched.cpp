// pt_b200/host/ptw/ptw_batched.cpp -- ptw_threads with batched move testing.
//
// Same command line, same inputs and the same three output tables as the
// reference's driver (/root/reference/src/ptw_threads.cpp:33-98); the one difference
// is the move tester: for --optimization-radius >= 0 the wanderers get a
// pt::move_tester::BatchedNeighborhood, which scores a tree's whole NNI
// neighbourhood in one engine call instead of one move at a time.  Everything else
// (option parsing, guru, wanderers, authority, tables) is pt's own code, linked
// unchanged.
#include <algorithm>
#include <cstdio>
#include <exception>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include <libpll/pll.h>
#include <model.hpp>
#include <pll_partition.hpp>
#include <pll_util.hpp>

#include "batched_neighborhood.hpp"
#include "visit_prefetcher.hpp"
#include "compressed_tree.hpp"
#include "guru.hpp"
#include "options.hpp"

#include <chrono>
#include <cstdlib>

namespace {

// one "<lnl>\t<newick>" line per table entry
void DumpTable(pt::TreeTable& table, const std::string& path)
{
  if (path.empty()) return;
  std::ofstream out(path);
  if (!out) throw std::invalid_argument("could not open " + path + " for writing");
  auto locked = table.lock_table();
  for (auto& entry : locked) out << entry.second << "\t" << entry.first << "\n";
}

struct TreeSet {
  std::vector<pll_utree_t*> trees;
  ~TreeSet()
  {
    for (pll_utree_t* t : trees) pll_utree_destroy(t, pt::pll::cb_erase_data);
  }
};

// wall-clock marks of the process's phases, printed with PTGPU_HOST_STATS
struct Phases {
  std::chrono::steady_clock::time_point last = std::chrono::steady_clock::now();
  bool on = std::getenv("PTGPU_HOST_STATS") != nullptr;
  void mark(const char* what)
  {
    const auto now = std::chrono::steady_clock::now();
    if (on) std::fprintf(stderr, "[ptw_batched] %-34s %8.3f s\n", what, std::chrono::duration<double>(now - last).count());
    last = now;
  }
};

int Run(int argc, const char* argv[])
{
  Phases phases;
  pt::Options options = pt::ParseArguments(argc, argv);

  std::shared_ptr<const pt::move_tester::BatchedNeighborhood> batched;
  if (options.optimization_radius >= 0) {
    batched = std::make_shared<pt::move_tester::BatchedNeighborhood>(options);
    options.move_tester = batched;
  }

  // visits of queued moves ahead of time (pt_b200/host/ptw/visit_prefetcher.hpp: bit-identical
  // to the in-place visit, tables line for line).  Worth it for few wanderers, whose visits would
  // otherwise run one after the other; PTGPU_PREFETCH_VISITS=n overrides, 0 turns it off.
  if (batched) {
    const size_t threads = std::max<size_t>(1, options.thread_count);
    pt::move_tester::VisitPrefetcher::Instance().Start(threads == 1 ? 6 : (threads <= 4 ? 2 : 0),
                                                       options.optimization_radius);
  }

  TreeSet input;
  input.trees = pt::pll::ParseMultiNewick(options.newick_path);
  if (input.trees.empty())
    throw std::invalid_argument("No valid trees found in " + options.newick_path);

  std::vector<std::string> labels, sequences;
  pt::pll::ParseFasta(options.fasta_path, input.trees[0]->tip_count, labels, sequences);
  pt::CompressedTree::BuildDictionary(labels);
  const pt::pll::Model model = pt::pll::ParseRaxmlInfo(options.raxml_path, options.rate_categories);

  phases.mark("arguments, trees, alignment");
  pt::Guru guru(options, input.trees, model, labels, sequences);
  phases.mark("guru (first partition, CUDA context)");
  guru.Start();
  guru.Wait(options.poll_ms);
  phases.mark("the walk");
  if (!options.skip_filtering) guru.FilterGoodTreeTable();

  DumpTable(guru.GetGoodTreeTable(), options.good_trees_path);
  DumpTable(guru.GetVisitedTreeTable(), options.visited_trees_path);
  DumpTable(guru.GetTestedTreeTable(), options.tested_trees_path);
  phases.mark("tables written");

  if (pt::move_tester::VisitPrefetcher::Instance().enabled()) {
    const auto visits = pt::pll::Partition::VisitStatistics();
    pt::move_tester::VisitPrefetcher& pf = pt::move_tester::VisitPrefetcher::Instance();
    std::fprintf(stderr, "visits ahead of time: %llu submitted (%llu declined), %llu taken by OptimizeAllBranches, "
                         "%llu computed in place; %llu neighbourhoods scored ahead, %llu visits chained after a failed "
                         "sibling, %llu queued by workers walking ahead\n",
                 pf.submitted(), pf.declined(), visits.first, visits.second, pf.neighborhoods_taken(),
                 pf.chained(), pf.walked_ahead());
  }
  pt::move_tester::VisitPrefetcher::Instance().Shutdown();   // before the partitions and the CUDA runtime go
  phases.mark("workers and their engines released");
  if (batched) {
    const auto stats = batched->Statistics();
    std::fprintf(stderr, "batched move testing: %llu moves answered from %llu engine batches\n",
                 static_cast<unsigned long long>(stats.second),
                 static_cast<unsigned long long>(stats.first));
  }
  return 0;
}

} // namespace

int main(int argc, const char* argv[])
{
  try {
    return Run(argc, argv);
  } catch (const std::exception& e) {
    std::cerr << "ptw_batched: " << e.what() << "\n";
    return 1;
  }
}
