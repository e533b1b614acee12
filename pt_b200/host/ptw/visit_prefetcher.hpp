// pt_b200/host/ptw/visit_prefetcher.hpp -- the visits of queued moves, computed ahead of time
// (SURVEY 8f-1).
//
// Wanderer::QueueMoves (src/wanderer.cpp:242-285) queues every move that passes its test and
// Wanderer::MoveForward (:150-210) later visits each of them: apply the NNI, InvalidateEdgeClvs,
// clone, Partition::OptimizeAllBranches, LogLikelihood.  On a small alignment such a visit is a
// chain of several thousand dependent device phases (6-10 ms on 8 CTAs) and a one-wanderer walk
// spends most of its time waiting for them one after the other.  BatchedNeighborhood hands a move
// to this class the moment it passes its test; worker threads, each with a Replica() of the
// wanderer's partition (own engine + stream on the same GPU), compute the visit and register the
// optimised lengths with pt::pll::Partition::PublishVisit; when MoveForward gets to the move, its
// OptimizeAllBranches finds them there.  The wanderer still commits visits in its own
// depth-first order.
//
// LINE FOR LINE, and how.  The reference's visit is not a function of the tree alone: it starts
// from the PARENT tree's CLVs with only the two ends of the NNI edge flagged invalid
// (PTPLL/pll_util.cpp:353-366), so the CLVs between that edge and the root are stale when the
// first branches are optimised, and the optimum found (to OptimizeAllBranches' own 1e-6
// tolerance) depends on it in the last digits.  A worker therefore reproduces that very start
// state -- FULL traversal of the parent tree T at the node the tree is oriented at when the visit
// starts, then the NNI and the two invalid flags -- and runs the identical procedure: same
// kernels, same inputs, same bits.  That node is T's virtual root when the wanderer returns to T
// by MoveBack (src/wanderer.cpp:232-239); for the FIRST move of T's queue it is the node the move
// tester tested last (the reference's testers leave the tree oriented there; the batched tester
// registers it with Partition::DeferTesterOrientation, the partition produces the state when the
// visit starts and hands the moves noted so far to Flush).  The result is registered under a key that names the start state (content, tree,
// flags: Partition::VisitKey); a visit whose actual start state is another one -- the move
// after a visit that did NOT lead to a good tree starts from that tree's CLVs instead -- finds
// nothing under its key and is computed in place, as before.  tests/test_gpu_dropin.py requires
// the good / visited / tested tables to be identical to the reference's, scores included.
//
// The same workers also score the NNI neighbourhood of every tree they have just optimised
// (Partition::ScoreNniNeighborhood is a pure function of the tree, its lengths and the model:
// it recomputes all matrices and directional CLVs itself) and register the scores under the
// optimised tree's fingerprint; BatchedNeighborhood takes them from there when the wanderer,
// having found the tree good, asks about its first move.
//
// Sibling chains.  The one start state that is NOT "exact CLVs of the parent" is the visit of a
// move whose preceding sibling did not lead to a good tree: the partition then still holds THAT
// tree's CLVs (src/wanderer.cpp:176-186), and the parent tree's flags carry the invalidated edge
// ends of every sibling tried since its last FULL traversal.  The workers follow each queue the way
// the wanderer will (Scan): move 0 from the exact parent; after a visit whose optimised tree is
// below the authority's threshold the next move from the state that visit left (the failed tree's
// exact CLVs -- still on the replica if the same worker goes on, otherwise put there by one FULL
// traversal of the kept optimised tree -- and the flags of the run of failures); after a good tree
// the next move from the exact parent again (MoveBack's FULL traversal, :232-239).  The threshold
// only ever rises (src/authority.cpp:205-211), so "failed" is final while "good" is not: every
// queue is followed again whenever the threshold has moved (the wanderer's next question shows
// it), and the visits a changed verdict calls for are started then -- the wanderer, walking depth
// first, returns to the later moves of a queue long after.  Both variants of a move may thus be
// computed; only the one whose start state occurs is ever taken.  Rejected trees get no
// neighbourhood scores: the wanderer never queues moves from them.
//
// Walking ahead (PTGPU_PREFETCH_DEPTH=d, default 0 = off).  A depth-first walk descends into a tree
// the moment it has been found good, so the visit of its first move is only asked for when the
// wanderer gets there.  With d > 0 a worker that has optimised a tree and scored its neighbourhood
// submits the visits of that tree's promising moves itself (tree good and move score above the
// authority's present threshold, move not leading to a tree already visited or already speculated
// on: what QueueMoves will decide, read from the same Authority), up to d levels ahead of the
// wanderer.  A guess that turns out wrong costs device time, never correctness: results are only
// ever taken by exact start state.  Measured on DS1 and the DS8-shape stand-in
// (profiles/r2_depth_sweep.log): no gain at any depth -- one level ahead starts a visit only
// ~0.5 ms before the wanderer would have asked for it, deeper levels spend as much device time on
// trees reached from another parent as they hide -- hence off by default.
//
// On by default in ptw_batched: Start(n) spawns n worker threads (PTGPU_PREFETCH_VISITS=n
// overrides, 0 = off).  It pays for few wanderers, whose visits otherwise run one after the
// other; with many wanderers the visits of different wanderers already overlap on the device.
#ifndef PT_B200_HOST_PTW_VISIT_PREFETCHER_HPP_
#define PT_B200_HOST_PTW_VISIT_PREFETCHER_HPP_

#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <thread>
#include <unordered_set>
#include <vector>

#include <libpll/pll.h>
#include <model.hpp>
#include <pll_partition.hpp>

#include "authority.hpp"

namespace pt { namespace move_tester {

class VisitPrefetcher
{
 public:
  static VisitPrefetcher& Instance();

  bool enabled() const { return !workers_.empty() && !disabled_.load(std::memory_order_relaxed); }

  // before the walk starts; PTGPU_PREFETCH_VISITS, if set, takes the place of `workers`.
  // `radius` is the move tester's (the workers score neighbourhoods with it).
  void Start(int workers, int radius);

  // The move (node, type) of the wanderer's tree `tree` has passed its test; `owner` is the
  // wanderer's partition, whose content is exact for `tree` (it has just been visited).
  // Nothing the wanderer owns is modified; returns at once.  Never throws: a failure turns
  // prefetching off for the rest of the run and the visit is computed in place.
  // The move is only NOTED: the start state of the queue's first visit depends on the node the
  // move tester tests LAST (the tree is left oriented at it: Partition::DeferTesterOrientation), so
  // the visits are queued by Flush, which the partition calls when that state has been produced.
  void Submit(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* node, int type,
              const Authority& authority) noexcept;
  // `tree` is now oriented at `at` (its edge ends invalid if `edge_invalid`), the partition exact for
  // it: queue the noted moves of `tree` -- the first from this state, the others from the exact
  // tree at its virtual root (where MoveBack leaves it); sibling chains follow as visits finish
  void Flush(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* at, bool edge_invalid) noexcept;

  // the scores of `tree`'s NNI neighbourhood if a worker has computed (or is computing) them:
  // node pointers are those of `tree`.  false = not registered (or the computation failed).
  bool TakeNeighborhood(pll_utree_t* tree, std::vector<pll::Partition::NniScore>* out);

  // joins the workers and destroys their engines (before the CUDA runtime goes away)
  void Shutdown();

  unsigned long long submitted() const { return submitted_; }
  unsigned long long declined() const { return declined_; }
  unsigned long long neighborhoods_taken() const { return hoods_taken_; }
  unsigned long long walked_ahead() const { return walked_ahead_; }
  unsigned long long chained() const { return chained_; }

  ~VisitPrefetcher();

 private:
  VisitPrefetcher();

  struct TreeHolder {
    pll_utree_t* t = nullptr;
    explicit TreeHolder(pll_utree_t* tree) : t(tree) {}
    TreeHolder(const TreeHolder&) = delete;
    TreeHolder& operator=(const TreeHolder&) = delete;
    ~TreeHolder();
  };
  // a finished visit: the optimised tree (lengths as OptimizeAllBranches left them) and its lnL
  struct Done {
    double lnl;
    std::shared_ptr<TreeHolder> tree;
  };
  typedef std::pair<size_t, size_t> VisitId;   // (index in the queue, first move of the run of failures it
                                               //  ends; == index for the visit from the exact parent)

  // the moves of one QueueMoves call that reached the prefetcher, in order, and the tree they
  // are moves of (a private clone carrying the flags a FULL traversal at its root leaves)
  struct Batch {
    unsigned long long id = 0;
    pll_utree_t* parent = nullptr;         // flags: FULL traversal at the virtual root
    pll_utree_t* parent_first = nullptr;   // flags: as the move tester left them (what the queue's first visit,
                                           // and a run of failures that starts with it, sees)
    int device = 0;
    pll::Model model;
    unsigned int parent_root_index = 0;
    std::vector<std::pair<unsigned int, int>> moves;   // (node_index, type), submission order
    std::unordered_set<CompressedTree> topologies;     // of the trees the moves lead to (duplicates are dropped)
    std::map<VisitId, Done> done;                      // finished visits of this queue
    std::map<VisitId, bool> verdict;                   // taken by the wanderer: was the tree good THEN (final)
    std::set<VisitId> started;                         // visits after a failed sibling already started
    ~Batch();
  };

  struct Task {
    std::shared_ptr<Batch> group;     // null for visits queued by workers walking ahead
    size_t index = 0, run_start = 0;  // position in group->moves; first move of the present run of failed siblings
    std::shared_ptr<TreeHolder> previous;   // the failed sibling's optimised tree this visit starts from (null: exact parent)
    int device;
    pll::Model model;
    pll_utree_t* tree = nullptr;      // private clone WITH the move applied and the start flags set
    unsigned int node_index;          // the move, by record index
    unsigned int parent_root_index;   // where the parent tree's exact CLVs are oriented when the visit starts
    bool parent_edge_invalid = false; // ... with the two ends of that node's edge invalid (radius-0 tester)
    int type;
    int depth = 0;                    // levels ahead of the wanderer (0: submitted by the wanderer)
    unsigned long long batch;         // moves queued by the same QueueMoves share a batch
    std::promise<pll::Partition::VisitResult> result;
    ~Task();
  };
  // a visit Scan found missing: move `index` of the queue after the failed visit `previous`
  struct Job {
    std::shared_ptr<Batch> group;
    size_t index, run_start;
    std::shared_ptr<TreeHolder> previous;
    std::vector<std::pair<unsigned int, int>> run;   // the failed moves since the last exact state
    std::pair<unsigned int, int> move;
  };
  struct IndexedScore { unsigned int node_index; int type; double length, lnl; };
  typedef std::vector<IndexedScore> Neighborhood;

  void Work();
  void SubmitOrThrow(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* node, int type);
  // queues the visit of move (node, type) of `tree`, whose exact CLVs (oriented at its virtual
  // root) the visiting partition will hold; false if that visit is already known
  // (`at` / `edge_invalid`: the orientation of the queue's first visit, null = the virtual root)
  bool Enqueue(int device, const pll::Model& model, pll_utree_t* tree, pll_unode_t* node, int type, int depth,
               const pll_unode_t* at = nullptr, bool edge_invalid = false);
  // one visit on `rep`; returns the optimised tree's lnL.  The start state: task.previous null --
  // the exact parent (made here); else the failed sibling's exact CLVs, which `rep` already holds
  // if `state_on_replica`, else made here by a FULL traversal of task.previous
  double RunVisit(Task& task, pll::Partition& rep, bool state_on_replica,
                  const std::function<void(double)>& before_release);
  // follows the queue as the wanderer will under the present threshold; the visits that calls for
  // and nobody has started yet (marked started here).  Takes mutex_.
  std::vector<Job> Scan(const std::shared_ptr<Batch>& group);
  std::unique_ptr<Task> BuildSibling(const Job& job);
  void NoteThreshold(const Authority& authority);
  // the wanderer has taken this visit's result (called on its thread, from OptimizeAllBranches)
  void Taken(const std::shared_ptr<Batch>& group, VisitId id, double lnl);
  void StartJobs(const std::shared_ptr<Batch>& group);
  void WalkAhead(const Task& done, pll::Partition& rep, const std::vector<pll::Partition::NniScore>& scored);

  // One pool of replicas per device, shared by every wanderer on it (all partitions of a ptw
  // run hold the same alignment).  Replicas are made on a wanderer's own thread (Submit): a
  // worker must not read an owner's state while its thread uses it.  At most as many per
  // device as there are workers and as the memory budget allows (PTGPU_PREFETCH_MAX_MB).
  struct Pool {
    std::vector<std::unique_ptr<pll::Partition>> free;
    size_t made = 0;
  };
  std::map<int, Pool> pools_;

  // moves that have passed their test, waiting for Flush: per owner, of one tree
  struct Noted {
    const pll_utree_t* tree = nullptr;
    std::vector<std::pair<unsigned int, int>> moves;   // (node_index, type)
  };
  std::map<const pll::Partition*, Noted> noted_;

  std::vector<std::thread> workers_;
  std::mutex mutex_;
  std::condition_variable wake_, replica_back_;
  std::deque<std::unique_ptr<Task>> queue_;
  bool stop_ = false;
  bool rescan_ = false;           // the threshold has moved: follow every queue again
  double last_threshold_ = 0.0;
  bool have_threshold_ = false;
  std::atomic<bool> disabled_{false};
  unsigned long long batch_ = 0;
  std::map<std::pair<const pll_utree_t*, uint64_t>, std::shared_ptr<Batch>> batches_;   // by (tree, its fingerprint)
  std::deque<std::pair<const pll_utree_t*, uint64_t>> batch_order_;
  std::atomic<unsigned long long> chained_{0};
  unsigned long long submitted_ = 0, declined_ = 0;

  std::map<uint64_t, std::shared_future<Neighborhood>> hoods_;
  std::deque<uint64_t> hood_order_;
  std::atomic<unsigned long long> hoods_taken_{0}, walked_ahead_{0};
  std::atomic<const Authority*> authority_{nullptr};
  std::unordered_set<CompressedTree> speculated_;   // topologies a worker has already walked ahead to
  int radius_ = 0;
  int max_depth_ = 0;         // PTGPU_PREFETCH_DEPTH
  bool score_hoods_ = true;   // PTGPU_PREFETCH_NEIGHBORHOODS=0 turns it off
};

} } // namespace pt::move_tester

#endif
