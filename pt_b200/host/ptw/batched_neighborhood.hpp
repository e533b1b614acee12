// pt_b200/host/ptw/batched_neighborhood.hpp -- a pt::MoveTester that scores ALL NNI
// candidates of a tree in one engine call and answers the wanderer's per-move
// questions from that result.
//
// The reference tests moves one at a time: Wanderer::QueueMoves
// (/root/reference/src/wanderer.cpp:242-285) calls MoveTester::EvaluateMove for up to
// 2(n-3) moves of the same tree, and each call to SingleBranchOptimizer /
// BranchNeighborhoodOptimizer (src/move_tester/*.cpp) costs several traversals and
// ~10-100 likelihood-kernel calls.  This tester plugs into the same interface
// (src/move_tester.hpp); on the first question about a tree it calls
// pt::pll::Partition::ScoreNniNeighborhood(tree, radius) -- one batched launch for
// radius 0, one persistent kernel for radius >= 1 -- and keeps the scores until the
// wanderer asks about a different tree.  Scores, and therefore the accepted-move
// sequence and the good / visited / tested tables, are those of the reference's
// testers (same arithmetic; tests/test_gpu_dropin.py compares the tables).
//
// Compiled against pt's own src/ headers (a maintainer would drop it into
// src/move_tester/ and select it in options.cpp).
#ifndef PT_B200_BATCHED_NEIGHBORHOOD_HPP_
#define PT_B200_BATCHED_NEIGHBORHOOD_HPP_

#include <cstdint>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include <libpll/pll.h>
#include <pll_partition.hpp>

#include "common.hpp"
#include "move_tester.hpp"
#include "options.hpp"

namespace pt { namespace move_tester {

class BatchedNeighborhood : public MoveTester {
 public:
  explicit BatchedNeighborhood(const Options& options);
  ~BatchedNeighborhood() override;

  std::pair<bool, double> EvaluateMove(pll::Partition& partition, pll_utree_t* tree,
                                       pll_unode_t* node, MoveType type,
                                       const Authority& authority) const override;

  // engine calls made / questions answered so far (for the run summary)
  std::pair<uint64_t, uint64_t> Statistics() const;

 private:
  struct Scores {
    const pll_utree_t* tree = nullptr;
    uint64_t fingerprint = 0;
    std::vector<pll::Partition::NniScore> moves;
  };

  static uint64_t Fingerprint(const pll_utree_t* tree);

  const int radius_;
  mutable std::mutex mutex_;                                  // guards the map, not the entries:
  mutable std::map<const pll::Partition*, Scores> by_partition_;  // an entry belongs to its wanderer's thread
  mutable uint64_t batches_ = 0, questions_ = 0, neighborhood_ns_ = 0;
};

} } // namespace pt::move_tester

#endif // PT_B200_BATCHED_NEIGHBORHOOD_HPP_
