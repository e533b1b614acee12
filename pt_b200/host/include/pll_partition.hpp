// pt_b200/host/include/pll_partition.hpp -- pt::pll::Partition over the B200
// engine (include/ptgpu.h).
//
// Drop-in for the reference's lib/libptpll/src/pll_partition.hpp:23-96: same
// class name, constructor, public methods, constants and exception behaviour,
// so src/wanderer.cpp, src/guru.cpp, the move testers and both Catch test files
// of the reference compile against it unchanged.  What differs is behind the
// methods: there is no pll_partition_t; every numeric step is a call into the
// extern "C" engine, which runs hand-written sm_100a kernels (no CPU path).
//
//   reference member (pll_partition.hpp)            here
//   PartitionPtr partition_ (pll_partition_t)       ptgpu_engine* engine_
//   double* sumtable_ (host scratch, :45)           engine-owned device sumtable
//   travbuffer_/branch_lengths_/matrix_indices_/    same host scratch (the traversal
//   operations_ (:39-42)                            descriptor is still built on the host)
//
// Threading is the reference's: one Partition per wanderer thread, methods not
// re-entrant per object.  Each Partition owns one engine = one CUDA stream on
// one device; devices are assigned round-robin over the visible GPUs
// (wanderer-parallel) unless PTGPU_DEVICE pins one.
//
// Site shards (BASELINE configs[2]): with PTGPU_SITE_SHARDS=n (n <= 8) a
// Partition holds n engines, each with a contiguous, tile-aligned range of the alignment's
// sites, on n consecutive devices (wrapping round when there are fewer).  The public surface is
// unchanged: every method issues the same engine call on all shards at once (one host thread
// per shard), the engines sum the per-tree scalars among themselves over NVLink peer memory
// (ptgpu_comm_init), and all shards return the same numbers.  OptimizeAllBranches is then ONE
// persistent kernel per GPU with no host round trip inside a pass.
#ifndef PT_B200_HOST_PLL_PARTITION_HPP_
#define PT_B200_HOST_PLL_PARTITION_HPP_

#include <cstdint>
#include <functional>
#include <future>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include <libpll/pll.h>

#include "model.hpp"
#include "pll_util.hpp"

struct ptgpu_engine;

namespace pt { namespace pll {

// MAX_ITER: Max iterations when optimizing branch lengths.
const unsigned int MAX_ITER = 32;
const double EPSILON = 1e-6; // Threshold for detecting zero.

class Partition
{
 private:
  const unsigned int tip_node_count_;
  bool map_mode_;
  ptgpu_engine* engine_;
  Model model_;
  unsigned int site_count_;

  // scratch buffers for TraversalUpdate()
  std::vector<pll_unode_t*> travbuffer_;
  std::vector<double> branch_lengths_;
  std::vector<unsigned int> matrix_indices_;
  std::vector<pll_operation_t> operations_;

 public:
  Partition(pll_utree_t* tree,
            const Model& model,
            const std::vector<std::string>& labels,
            const std::vector<std::string>& sequences,
            bool map_mode = false);

  Partition(Partition&& rhs);

  Partition(const Partition& rhs) = delete;
  Partition& operator=(const Partition& rhs) = delete;
  Partition& operator=(Partition&& rhs) = delete;

  ~Partition();

  double LogLikelihood(pll_unode_t* tree, double* per_site_lnl = nullptr);
  double LogMarginalLikelihood(pll_unode_t* root);

  unsigned int TraversalUpdate(pll_unode_t* root, TraversalType type);
  void UpdateBranchLength(pll_unode_t* node, double length);

  double OptimizeBranch(pll_unode_t* node);
  void OptimizeAllBranchesOnce(pll_unode_t* tree, bool use_pllmod = false);
  void OptimizeAllBranches(pll_unode_t* tree, bool use_pllmod = false);
  void OptimizeBranchNeighborhood(pll_unode_t* node, int radius);

  double OptimizeModelOnce(pll_unode_t* tree);
  double OptimizeModel(pll_unode_t* tree);

  double OptimizeAllBranchesAndModel(pll_unode_t* tree, bool use_pllmod = false);

  Model GetModel() const;
  void SetModel(const Model& model);

  unsigned int tip_node_count() const;
  unsigned int inner_node_count() const;
  unsigned int node_count() const;
  unsigned int branch_count() const;

  // ---- additions of the B200 build (not in the reference surface) ----

  // Scores every NNI candidate Wanderer::QueueMoves would test on `tree`
  // (src/wanderer.cpp:242-285: inner-node entries of tree->nodes in order,
  // pendant edges skipped, LEFT then RIGHT) in ONE batched engine call with
  // SingleBranchOptimizer semantics (src/move_tester/single_branch_optimizer.cpp:23-64):
  // optimised central branch length and log-likelihood per candidate.  The
  // tree, its lengths and the partition's CLV orientation flags are left
  // untouched (the batch works on a private set of directional CLVs).
  struct NniScore {
    pll_unode_t* node;   // the edge node passed to pll_utree_nni
    int type;            // PLL_UTREE_MOVE_NNI_LEFT / _RIGHT
    double length;       // optimised length of the central branch
    double lnl;          // log-likelihood at that length
  };
  // radius >= 1: BranchNeighborhoodOptimizer semantics instead
  // (src/move_tester/branch_neighborhood_optimizer.cpp:27-80: pll-modules' local
  // branch-length optimisation around the moved edge, then the log-likelihood);
  // `length` is then the optimised length of the central branch.
  std::vector<NniScore> ScoreNniNeighborhood(pll_utree_t* tree, int radius = 0);

  int device() const { return device_; }
  ptgpu_engine* engine() { return engine_; }
  // site shards of this partition (1 unless PTGPU_SITE_SHARDS asked for more)
  unsigned int shard_count() const { return static_cast<unsigned int>(shards_.size()); }

  // ---- visits computed ahead of time (pt_b200/host/ptw/visit_prefetcher) ----
  //
  // Wanderer::MoveForward visits every move it queued: apply the NNI, InvalidateEdgeClvs, clone,
  // OptimizeAllBranches, LogLikelihood (src/wanderer.cpp:150-210).  What that OptimizeAllBranches
  // computes is a function of three things only: the partition's CLV / P-matrix CONTENT when it
  // starts, the tree (topology + lengths) and the tree's per-record orientation flags -- the
  // reference starts it from the PARENT tree's CLVs with just the two ends of the NNI edge
  // flagged invalid, so the CLVs between that edge and the root are stale on purpose, and the
  // result depends on it.  The content is known exactly at the two moments that matter: after a
  // FULL TraversalUpdate and after an OptimizeAllBranches it is "the exact CLVs of that tree at
  // those lengths, oriented at that root" (every stale CLV has been recomputed by then).  So a
  // visit can be computed ahead of time on another engine that is put into the same state (FULL
  // traversal of the parent tree, same flags), and is then bit-identical to the one the
  // wanderer would have run: VisitKey() names the start state, PublishVisit() registers a
  // pending result under it, and OptimizeAllBranches() of a non-replica partition takes a
  // registered result whose key matches its own present state instead of computing -- it writes
  // the optimised lengths into the tree and brings the CLVs up with one FULL traversal, which
  // is the state the computation would have left.  Any other start state simply finds nothing
  // registered and is computed in place.
  struct VisitResult {
    std::vector<double> lengths;   // by node_index of every record of the tree
    std::function<void()> taken;   // if set: called (on the taker's thread) when the result is taken
  };
  // `directional`: with the buffers ScoreNniNeighborhood needs (4x the CLV memory)
  std::unique_ptr<Partition> Replica(bool directional = false) const;
  bool is_replica() const { return is_replica_; }
  // order-independent hash of (node_index, back->node_index, length) over all records
  static uint64_t TreeFingerprint(const pll_utree_t* tree);
  // "the exact CLVs of the tree with this fingerprint, oriented at record `root_index`"
  static uint64_t ExactContentId(uint64_t tree_fingerprint, unsigned int root_index);
  // key of OptimizeAllBranches(root) started from content `content_id` on the tree `root`
  // belongs to, flags included (0 if the tree cannot be traversed)
  static uint64_t VisitKeyFor(uint64_t content_id, pll_unode_t* root, unsigned int node_count);
  // the same for this partition's present content; 0 when that is not known exactly
  uint64_t VisitKey(pll_unode_t* root) const;
  // the present content's id (ExactContentId of the tree and root it is exact for); 0 = not known
  uint64_t ContentId() const { return content_id_; }

  // The state the reference's move testers leave behind.  BranchNeighborhoodOptimizer ends every test
  // with TraversalUpdate(tested node, FULL) on the wanderer's tree
  // (src/move_tester/branch_neighborhood_optimizer.cpp:71-73); SingleBranchOptimizer leaves the tree
  // oriented at the tested node with the two ends of its edge invalid (single_branch_optimizer.cpp:33-60).
  // After QueueMoves the tree is therefore oriented at the LAST node tested, and the visit that follows
  // (src/wanderer.cpp:150-190) starts from those validity flags -- which decide which stale CLVs its
  // first branches reuse.  A tester that scores the neighbourhood in one call and leaves the partition
  // alone registers that state here after every test; it is produced -- one FULL traversal at `tested`
  // on `tree`, the edge invalidated if asked, the invalidations this thread has made since replayed
  // (pll_util.hpp InvalidationLog), the flags copied onto the clone being visited -- when the next
  // OptimizeAllBranches starts, and forgotten if a FULL traversal comes first (MoveBack,
  // src/wanderer.cpp:232-239).  `tree` must still be alive then: it is, the wanderer visits clones of it.
  void DeferTesterOrientation(pll_utree_t* tree, pll_unode_t* tested, bool invalidate_edge);
  // called (on the visiting thread) right after a deferred orientation has been produced, before the
  // visit looks for a registered result: (partition, tree, node it is now oriented at, edge invalidated)
  static void SetOrientationHook(std::function<void(Partition&, pll_utree_t*, pll_unode_t*, bool)> hook);
  // (`bytes`: about how much the result will hold, for the registry's memory budget)
  static void PublishVisit(uint64_t key, std::shared_future<VisitResult> result, size_t bytes = 0);
  static bool VisitKnown(uint64_t key);
  static void ClearVisits();
  // (visits served from the registry, visits computed here) since process start
  static std::pair<unsigned long long, unsigned long long> VisitStatistics();
  // device memory one Replica() of this partition takes, roughly
  size_t ReplicaBytes(bool directional = false) const;

 private:
  int device_;
  unsigned int directional_base_;  // first extra CLV / scaler buffer (0 = none allocated)
  bool stepwise_all_branches_ = false;  // PTGPU_STEPWISE_ALL_BRANCHES: one engine call per step (A/B testing)
  bool is_replica_ = false;
  // what the per-node CLVs and the P matrices hold, when that is known exactly
  // (ExactContentId); 0 = unknown
  void ProduceTesterOrientation(pll_unode_t* visit_root);
  pll_utree_t* pending_tree_ = nullptr;      // DeferTesterOrientation
  pll_unode_t* pending_node_ = nullptr;
  bool pending_invalidate_ = false;
  uint64_t content_id_ = 0;
  uint64_t content_fp_ = 0;   // TreeFingerprint of the tree content_id_ was formed from

  // all engines of this partition; shards_[0] == engine_, shard s holds the sites
  // [shard_begin_[s], shard_begin_[s + 1]) of the alignment
  std::vector<ptgpu_engine*> shards_;
  std::vector<unsigned int> shard_begin_;
  class ShardPool;
  std::shared_ptr<ShardPool> pool_;
  // runs call(engine, shard) for every shard (concurrently when there are several: a shard's
  // kernel waits for its peers') and throws what Check would for the first failure
  void OnShards(const std::function<int(ptgpu_engine*, unsigned int)>& call) const;
  void CreateEngines(unsigned int clv_buffers, unsigned int scale_buffers, unsigned int rate_cats,
                     unsigned int wanted_shards);
  void DestroyEngines();

  struct ReplicaTag {};
  Partition(const Partition& proto, ReplicaTag, bool directional);
  uint64_t FingerprintFrom(pll_unode_t* root);
  bool TakePublishedVisit(uint64_t key, pll_unode_t* root);

  void SetTipStates(pll_utree_t* tree, const std::vector<std::string>& labels,
                    const std::vector<std::string>& sequences);
  void AllocateScratchBuffers();
  void Check(int rc) const;
  double MapLogPrior(pll_unode_t* tree);
  void OptimizeAllBranchesFused(pll_unode_t* tree);
  unsigned int DirectionalClv(const pll_unode_t* record) const;
  int DirectionalScaler(const pll_unode_t* record) const;
  void UpdateDirectionalClvs(pll_utree_t* tree, const std::vector<unsigned int>& needed);
  std::vector<NniScore> ScoreNniNeighborhoodLocal(pll_utree_t* tree, int radius);
  double LocalBranchOptimization(pll_unode_t* tree, double bl_min, double bl_max,
                                 double tolerance, int smoothings, int radius);
  bool OptimizeEdgeBounded(pll_unode_t* edge, double bl_min, double bl_max,
                           double tolerance, int radius);
};

inline unsigned int Partition::tip_node_count() const { return tip_node_count_; }
inline unsigned int Partition::inner_node_count() const { return tip_node_count_ - 2; }
inline unsigned int Partition::node_count() const { return tip_node_count() + inner_node_count(); }
inline unsigned int Partition::branch_count() const { return node_count() - 1; }

} } // namespace pt::pll

#endif // PT_B200_HOST_PLL_PARTITION_HPP_
