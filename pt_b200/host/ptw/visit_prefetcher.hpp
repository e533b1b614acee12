// pt_b200/host/ptw/visit_prefetcher.hpp -- the visits of queued moves, computed in the
// background.
//
// Wanderer::QueueMoves (src/wanderer.cpp:242-285) queues every move that passes its test and
// Wanderer::MoveForward (:150-210) later visits each of them: clone the tree with the move
// applied, Partition::OptimizeAllBranches, LogLikelihood.  On a small alignment such a visit
// is a chain of several thousand dependent device phases (6-10 ms on 8 CTAs) and the walk
// spends most of its time waiting for them one after the other.  The visit of a queued move
// depends on nothing but the tree it leads to, so BatchedNeighborhood hands that tree to this
// class the moment the move passes its test; worker threads, each with a Replica() of the
// wanderer's partition (own engine + stream on the same GPU), run the identical call
// sequence on a private clone and register the optimised lengths with
// pt::pll::Partition::PublishVisit.  When MoveForward gets to the move, its
// OptimizeAllBranches finds the lengths there.  The wanderer still commits visits in its own
// depth-first order.
//
// OPT-IN, and why.  A visit computed here starts from CLVs that are all exact for the tree
// (FULL traversal).  The reference's own visit does not: Wanderer::MoveForward invalidates only
// the two CLVs at the ends of the NNI edge (InvalidateEdgeClvs, PTPLL/pll_util.cpp:353-366;
// src/wanderer.cpp:160-186), so CLVs between that edge and the virtual root stay flagged valid
// although the move changed what they summarise -- and after a move that did not lead to a
// good tree the partition still holds THAT tree's CLVs (src/wanderer.cpp:176-186) -- and the
// first branch optimisations of the visit read them.  OptimizeAllBranches converges to the
// same optimum either way, but only to its own tolerance (passes stop at |dlnL| < 1e-6), so
// optimised lengths differ in their last bits.  The in-place visit (the default) reproduces the
// reference's state and therefore its tables line for line; with prefetching, good and visited
// tables came out identical on DS1 and the three stand-ins, while a few of the ~10^4 tested-tree
// scores differ in the last of their 6 printed digits (tests/test_gpu_dropin.py).
//
// Start(n) spawns n worker threads (0 = off); PTGPU_PREFETCH_VISITS=n overrides.  It pays for ONE
// wanderer, whose visits otherwise run one after the other (DS8 stand-in: 27.6 -> 12.0 s with 8
// workers, DS1: 4.9 -> 3.6 s with 2); with many wanderers the visits of different wanderers
// already overlap on the device and more engines only add contention (measured).
#ifndef PT_B200_HOST_PTW_VISIT_PREFETCHER_HPP_
#define PT_B200_HOST_PTW_VISIT_PREFETCHER_HPP_

#include <condition_variable>
#include <deque>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include <libpll/pll.h>
#include <pll_partition.hpp>

namespace pt { namespace move_tester {

class VisitPrefetcher
{
 public:
  static VisitPrefetcher& Instance();

  bool enabled() const { return !workers_.empty(); }

  // before the walk starts; PTGPU_PREFETCH_VISITS, if set, takes the place of `workers`
  void Start(int workers);

  // `tree` is the wanderer's tree WITH the move applied (the caller undoes the move after
  // this returns); `owner` the wanderer's partition.  Returns at once.
  void Submit(pll::Partition& owner, pll_utree_t* tree);

  // joins the workers and destroys their engines (before the CUDA runtime goes away)
  void Shutdown();

  unsigned long long submitted() const { return submitted_; }

  ~VisitPrefetcher();

 private:
  VisitPrefetcher();

  struct Task {
    pll::Partition* owner;
    pll_utree_t* tree;   // private clone, destroyed by the worker
    unsigned long long batch;   // moves queued by the same QueueMoves share a batch
    std::promise<pll::Partition::VisitResult> result;
  };

  void Work();

  // Replicas are made on the wanderer's own thread (Submit), one per worker, the first time a
  // partition submits: a worker must not read the owner's state while its thread uses it.
  std::map<pll::Partition*, std::vector<std::unique_ptr<pll::Partition>>> free_replicas_;

  std::vector<std::thread> workers_;
  std::mutex mutex_;
  std::condition_variable wake_;
  std::deque<std::unique_ptr<Task>> queue_;
  bool stop_ = false;
  unsigned long long batch_ = 0;
  std::map<pll::Partition*, const pll_utree_t*> last_tree_;   // per owner: the tree of its latest batch
  unsigned long long submitted_ = 0;
};

} } // namespace pt::move_tester

#endif
