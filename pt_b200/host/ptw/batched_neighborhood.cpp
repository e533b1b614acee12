// pt_b200/host/ptw/batched_neighborhood.cpp -- see the header.
#include "batched_neighborhood.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

#include "authority.hpp"
#include "visit_prefetcher.hpp"

namespace pt { namespace move_tester {

BatchedNeighborhood::BatchedNeighborhood(const Options& options)
    : MoveTester(options), radius_(options.optimization_radius)
{
  if (radius_ < 0) throw std::invalid_argument("optimization radius must not be negative");
  if (options.marginal_mode)
    throw std::invalid_argument("marginal mode is not available with batched move testing");
  // once the partition has produced the state the reference's tester would have left (below), the
  // visits of the moves that passed their test are queued from exactly that state
  pll::Partition::SetOrientationHook([](pll::Partition& partition, pll_utree_t* tree, pll_unode_t* at, bool edge_invalid) {
    VisitPrefetcher::Instance().Flush(partition, tree, at, edge_invalid);
  });
}

BatchedNeighborhood::~BatchedNeighborhood()
{
  if (neighborhood_ns_)
    std::fprintf(stderr, "[ptgpu host] %-28s %10llu calls %10.3f s\n", "neighbourhood (take or score)",
                 static_cast<unsigned long long>(batches_), neighborhood_ns_ * 1e-9);
}

// Topology + branch lengths of the tree as the wanderer holds it: who each record
// is linked to and at which length (FNV-1a over the node array).  The wanderer
// edits trees in place between QueueMoves calls (clone / NNI / optimise), so the
// pointer alone does not identify what was scored.
uint64_t BatchedNeighborhood::Fingerprint(const pll_utree_t* tree)
{
  uint64_t h = 1469598103934665603ull;
  auto mix = [&h](uint64_t v) {
    for (int k = 0; k < 8; ++k) {
      h ^= (v >> (8 * k)) & 0xffu;
      h *= 1099511628211ull;
    }
  };
  const unsigned int n = tree->tip_count + tree->inner_count;
  for (unsigned int i = 0; i < n; ++i) {
    const pll_unode_t* r = tree->nodes[i];
    const pll_unode_t* first = r;
    do {
      uint64_t bits;
      std::memcpy(&bits, &r->length, sizeof bits);
      mix(r->node_index);
      mix(r->back->node_index);
      mix(bits);
      r = r->next;
    } while (r && r != first);
  }
  return h;
}

std::pair<bool, double> BatchedNeighborhood::EvaluateMove(pll::Partition& partition,
                                                          pll_utree_t* tree, pll_unode_t* node,
                                                          MoveType type,
                                                          const Authority& authority) const
{
  Scores* scores;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    scores = &by_partition_[&partition];
    ++questions_;
  }
  const uint64_t fp = Fingerprint(tree);
  static const bool stats = std::getenv("PTGPU_HOST_STATS") != nullptr;
  if (scores->tree != tree || scores->fingerprint != fp) {
    const auto t0 = std::chrono::steady_clock::now();
    // first question about this tree: score its whole NNI neighbourhood.  The
    // partition's per-node CLVs, the tree and its orientation flags stay as they
    // are (the batch works on directional and private buffers), which is the
    // state the reference's testers restore before they return.
    // (a background worker may already have scored this very tree: visit_prefetcher.hpp)
    if (!VisitPrefetcher::Instance().TakeNeighborhood(tree, &scores->moves))
      scores->moves = partition.ScoreNniNeighborhood(tree, radius_);
    scores->tree = tree;
    scores->fingerprint = fp;
    std::lock_guard<std::mutex> lock(mutex_);
    ++batches_;
    if (stats) neighborhood_ns_ += std::chrono::duration_cast<std::chrono::nanoseconds>(
                                       std::chrono::steady_clock::now() - t0).count();
  }
  for (const pll::Partition::NniScore& m : scores->moves) {
    if (m.node == node && m.type == static_cast<int>(type)) {
      // testing only: the score is not reported to the authority here
      const bool accept = m.lnl >= authority.GetThresholdScore();
      // the wanderer will visit this move if the authority grants it (src/wanderer.cpp:262-272):
      // start that visit now, on another engine (the wanderer's tree is not touched; never throws)
      if (accept) VisitPrefetcher::Instance().Submit(partition, tree, node, static_cast<int>(type), authority);
      // The reference's testers leave the wanderer's tree oriented at the node of the move they have
      // just tested (BranchNeighborhoodOptimizer: FULL traversal there; SingleBranchOptimizer: oriented
      // there, the edge's two ends invalid): the next visit starts from that.  Nothing is touched here;
      // the partition produces the state when -- if -- a visit follows (Partition::DeferTesterOrientation).
      partition.DeferTesterOrientation(tree, node, radius_ == 0);
      return std::make_pair(accept, m.lnl);
    }
  }
  throw std::logic_error("move is not an NNI across an inner edge of this tree");
}

std::pair<uint64_t, uint64_t> BatchedNeighborhood::Statistics() const
{
  std::lock_guard<std::mutex> lock(mutex_);
  return std::make_pair(batches_, questions_);
}

} } // namespace pt::move_tester
