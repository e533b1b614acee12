// pt_b200/host/ptw/visit_prefetcher.cpp -- see the header.
#include "visit_prefetcher.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <stdexcept>

#include <pll_util.hpp>

#include "ordered_tree.hpp"   // GetVirtualRoot

namespace pt { namespace move_tester {

namespace {

pll_unode_t* RecordByIndex(pll_utree_t* tree, unsigned int node_index)
{
  const unsigned int n = tree->tip_count + tree->inner_count;
  for (unsigned int i = 0; i < n; ++i) {
    pll_unode_t* first = tree->nodes[i];
    pll_unode_t* r = first;
    do {
      if (r->node_index == node_index) return r;
      r = r->next;
    } while (r && r != first);
  }
  throw std::logic_error("visit prefetcher: no record with that node index");
}

bool SameModel(const pll::Model& a, const pll::Model& b)
{
  return a.model_name == b.model_name && a.frequencies == b.frequencies &&
         a.subst_params == b.subst_params && a.category_rates == b.category_rates;
}

} // namespace

VisitPrefetcher::Task::~Task()
{
  if (tree) pll_utree_destroy(tree, pll::cb_erase_data);
}

VisitPrefetcher& VisitPrefetcher::Instance()
{
  static VisitPrefetcher instance;
  return instance;
}

VisitPrefetcher::VisitPrefetcher() {}

void VisitPrefetcher::Start(int workers, int radius)
{
  if (!workers_.empty()) return;
  int n = workers;
  radius_ = radius;
  if (const char* env = std::getenv("PTGPU_PREFETCH_VISITS")) n = std::atoi(env);
  if (const char* env = std::getenv("PTGPU_PREFETCH_NEIGHBORHOODS")) score_hoods_ = std::atoi(env) != 0;
  if (const char* env = std::getenv("PTGPU_PREFETCH_DEPTH")) max_depth_ = std::atoi(env);
  {
    std::lock_guard<std::mutex> lock(mutex_);
    stop_ = false;
  }
  disabled_ = false;
  for (int i = 0; i < n; ++i) workers_.emplace_back([this] { Work(); });
}

VisitPrefetcher::~VisitPrefetcher() { Shutdown(); }

void VisitPrefetcher::Shutdown()
{
  {
    std::lock_guard<std::mutex> lock(mutex_);
    stop_ = true;
  }
  wake_.notify_all();
  replica_back_.notify_all();
  for (std::thread& t : workers_)
    if (t.joinable()) t.join();
  workers_.clear();
  std::lock_guard<std::mutex> lock(mutex_);
  pools_.clear();   // the engines go here, while the CUDA runtime is still up
  queue_.clear();
  hoods_.clear();
  hood_order_.clear();
  speculated_.clear();
  last_tree_.clear();
  pll::Partition::ClearVisits();
}

void VisitPrefetcher::Submit(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* node, int type,
                             const Authority& authority) noexcept
{
  if (!enabled()) return;
  authority_.store(&authority, std::memory_order_relaxed);
  try {
    SubmitOrThrow(owner, tree, node, type);
  } catch (const std::exception& e) {
    // an optional optimisation must never disturb the walk: nothing of the wanderer's has been
    // touched (the move is applied to a private clone), so just stop prefetching
    disabled_ = true;
    std::fprintf(stderr, "visit prefetcher: turned off (%s)\n", e.what());
  } catch (...) {
    disabled_ = true;
  }
}

void VisitPrefetcher::SubmitOrThrow(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* node, int type)
{
  if (owner.is_replica() || owner.shard_count() > 1) return;
  const int device = owner.device();
  // a replica to run on: made here, on the owner's thread, within the memory budget
  size_t budget_mb = 8192;
  if (const char* env = std::getenv("PTGPU_PREFETCH_MAX_MB")) budget_mb = std::strtoull(env, nullptr, 10);
  const size_t per = std::max<size_t>(1, owner.ReplicaBytes(score_hoods_));
  const size_t allowed = std::min<size_t>(workers_.size(), (budget_mb << 20) / per);
  size_t made;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    made = pools_[device].made;
  }
  if (made < allowed) {
    std::unique_ptr<pll::Partition> rep = owner.Replica(score_hoods_);
    std::lock_guard<std::mutex> lock(mutex_);
    Pool& pool = pools_[device];
    pool.free.push_back(std::move(rep));
    pool.made += 1;
    made = pool.made;
    replica_back_.notify_one();
  }
  if (made == 0) {
    ++declined_;
    return;
  }
  Enqueue(device, owner.GetModel(), tree, node, type, 0);
}

bool VisitPrefetcher::Enqueue(int device, const pll::Model& model, pll_utree_t* tree, pll_unode_t* node, int type,
                              int depth)
{
  const unsigned int n_nodes = tree->tip_count + tree->inner_count;
  // The start state of the visit as Wanderer::MoveForward will produce it (src/wanderer.cpp:160-166)
  // when the partition holds the exact CLVs of `tree` oriented at its virtual root: the tree's
  // flags are then those a FULL traversal leaves, plus the two invalid ends of the NNI edge.
  // The clone is made the way the wanderer makes it -- NNI on (a copy of) its tree, THEN the clone
  // that is visited -- because pll_utree_clone rebuilds the node array from the topology it
  // finds: the candidates of the visited tree's neighbourhood are enumerated in that order.
  std::unique_ptr<Task> task(new Task);
  {
    struct Holder {
      pll_utree_t* t;
      ~Holder() { if (t) pll_utree_destroy(t, pll::cb_erase_data); }
    } parent{pll_utree_clone(tree)};
    if (!parent.t) throw std::runtime_error("pll_utree_clone failed");
    pll_utree_every(parent.t, pll::cb_copy_clv_traversal);
    std::vector<pll_unode_t*> buffer(n_nodes, nullptr);
    unsigned int count = 0;
    if (!pll_utree_traverse(GetVirtualRoot(parent.t), PLL_TREE_TRAVERSE_POSTORDER, pll::cb_full_traversal,
                            buffer.data(), &count))
      throw std::runtime_error("the tree's virtual root is not an inner node");
    pll_unode_t* moved = RecordByIndex(parent.t, node->node_index);
    pll_utree_nni(moved, type, nullptr);
    pll::InvalidateEdgeClvs(moved);
    task->tree = pll_utree_clone(parent.t);
    if (!task->tree) throw std::runtime_error("pll_utree_clone failed");
    pll_utree_every(task->tree, pll::cb_copy_clv_traversal);
  }
  task->parent_root_index = GetVirtualRoot(tree)->node_index;
  const uint64_t content = pll::Partition::ExactContentId(pll::Partition::TreeFingerprint(tree),
                                                          GetVirtualRoot(tree)->node_index);
  const uint64_t key = pll::Partition::VisitKeyFor(content, GetVirtualRoot(task->tree), n_nodes);
  if (!key || pll::Partition::VisitKnown(key)) return false;
  task->device = device;
  task->model = model;
  task->depth = depth;
  task->node_index = node->node_index;
  task->type = type;
  pll::Partition::PublishVisit(key, task->result.get_future().share());
  {
    std::lock_guard<std::mutex> lock(mutex_);
    std::pair<const pll_utree_t*, uint64_t>& last = last_tree_[device];
    if (last.first != tree || last.second != content) {
      last = std::make_pair(static_cast<const pll_utree_t*>(tree), content);
      ++batch_;
    }
    task->batch = batch_;
    queue_.push_back(std::move(task));
    ++submitted_;
  }
  wake_.notify_one();
  return true;
}

// The wanderer will find this tree good and test its moves (src/wanderer.cpp:199-205, 242-285):
// queue the visits its QueueMoves is going to queue, now.
void VisitPrefetcher::WalkAhead(const Task& done, pll::Partition& rep,
                                const std::vector<pll::Partition::NniScore>& scored)
{
  const Authority* authority = authority_.load(std::memory_order_relaxed);
  if (!authority || done.depth >= max_depth_) return;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    if (queue_.size() > 8 * workers_.size()) return;   // plenty to do already
  }
  const double threshold = authority->GetThresholdScore();
  if (rep.LogLikelihood(GetVirtualRoot(done.tree)) < threshold) return;   // not a good tree: no QueueMoves
  for (const pll::Partition::NniScore& m : scored) {
    if (m.lnl < threshold) continue;
    // Authority::ProposeMove only reads its tables (and applies + undoes the move on OUR clone)
    if (!const_cast<Authority*>(authority)->ProposeMove(done.tree, m.node, static_cast<MoveType>(m.type))) continue;
    {
      // a tree within reach of two good trees is visited from whichever the wanderer meets first:
      // speculate on it once only (from the first parent a worker sees)
      pll_utree_nni(m.node, m.type, nullptr);
      const CompressedTree key = authority->GetKey(done.tree);
      pll_utree_nni(m.node, m.type, nullptr);
      std::lock_guard<std::mutex> lock(mutex_);
      if (!speculated_.insert(key).second) continue;
    }
    if (Enqueue(done.device, done.model, done.tree, m.node, m.type, done.depth + 1)) ++walked_ahead_;
  }
}

bool VisitPrefetcher::TakeNeighborhood(pll_utree_t* tree, std::vector<pll::Partition::NniScore>* out)
{
  if (!enabled() || !score_hoods_) return false;
  const uint64_t fp = pll::Partition::TreeFingerprint(tree);
  std::shared_future<Neighborhood> pending;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    auto it = hoods_.find(fp);
    if (it == hoods_.end()) return false;
    pending = it->second;
    hoods_.erase(it);
  }
  try {
    const Neighborhood& scored = pending.get();
    std::vector<pll_unode_t*> by_index(tree->tip_count + 3 * tree->inner_count, nullptr);
    const unsigned int n = tree->tip_count + tree->inner_count;
    for (unsigned int i = 0; i < n; ++i) {
      pll_unode_t* first = tree->nodes[i];
      pll_unode_t* r = first;
      do {
        if (r->node_index < by_index.size()) by_index[r->node_index] = r;
        r = r->next;
      } while (r && r != first);
    }
    out->clear();
    for (const IndexedScore& m : scored) {
      if (m.node_index >= by_index.size() || !by_index[m.node_index]) return false;
      out->push_back(pll::Partition::NniScore{by_index[m.node_index], m.type, m.length, m.lnl});
    }
  } catch (...) {
    return false;   // computed in place instead, where the caller sees the error
  }
  ++hoods_taken_;
  return true;
}

void VisitPrefetcher::Work()
{
  for (;;) {
    std::unique_ptr<Task> task;
    std::unique_ptr<pll::Partition> rep;
    {
      std::unique_lock<std::mutex> lock(mutex_);
      wake_.wait(lock, [this] { return stop_ || !queue_.empty(); });
      if (stop_) break;
      // The wanderer walks depth first: the queue it fills last is the one it empties first
      // (move_queues_ is a stack, src/wanderer.cpp:122-131), and each queue from its front
      // (:157-158).  So: what the wanderer itself asked for before speculation, newest batch
      // first, oldest submission within it.
      size_t pick = 0;
      for (size_t i = 1; i < queue_.size(); ++i) {
        const Task& a = *queue_[i];
        const Task& b = *queue_[pick];
        if (a.depth < b.depth || (a.depth == b.depth && a.batch > b.batch)) pick = i;
      }
      task = std::move(queue_[pick]);
      queue_.erase(queue_.begin() + pick);
      Pool& pool = pools_[task->device];
      replica_back_.wait(lock, [&] { return stop_ || !pool.free.empty(); });
      if (stop_) break;   // the task's promise breaks: the wanderer computes the visit in place
      rep = std::move(pool.free.back());
      pool.free.pop_back();
    }
    try {
      if (!SameModel(rep->GetModel(), task->model)) rep->SetModel(task->model);
      pll_unode_t* moved = RecordByIndex(task->tree, task->node_index);
      // back to the parent tree: its exact CLVs, oriented at its root, are what the wanderer's
      // partition holds when it starts this visit
      pll_utree_nni(moved, task->type, nullptr);
      rep->TraversalUpdate(RecordByIndex(task->tree, task->parent_root_index), pll::TraversalType::FULL);
      // the move, as MoveForward applies it
      pll_utree_nni(moved, task->type, nullptr);
      pll::InvalidateEdgeClvs(moved);
      rep->OptimizeAllBranches(GetVirtualRoot(task->tree));
      pll::Partition::VisitResult res;
      const unsigned int n = task->tree->tip_count + task->tree->inner_count;
      res.lengths.assign(task->tree->tip_count + 3 * task->tree->inner_count, 0.0);
      for (unsigned int i = 0; i < n; ++i) {
        pll_unode_t* first = task->tree->nodes[i];
        pll_unode_t* r = first;
        do {
          if (r->node_index < res.lengths.size()) res.lengths[r->node_index] = r->length;
          r = r->next;
        } while (r && r != first);
      }
      // the neighbourhood of the optimised tree, for the wanderer's QueueMoves: announced before
      // the visit's result is released, so the wanderer cannot get ahead of the announcement
      std::promise<Neighborhood> hood;
      const bool score = score_hoods_ && radius_ >= 0;
      if (score) {
        const uint64_t fp = pll::Partition::TreeFingerprint(task->tree);
        std::lock_guard<std::mutex> lock(mutex_);
        if (hoods_.emplace(fp, hood.get_future().share()).second) hood_order_.push_back(fp);
        while (hood_order_.size() > 4096) {
          hoods_.erase(hood_order_.front());
          hood_order_.pop_front();
        }
      }
      task->result.set_value(std::move(res));
      if (score) {
        std::vector<pll::Partition::NniScore> moves;
        try {
          moves = rep->ScoreNniNeighborhood(task->tree, radius_);
          Neighborhood scored;
          for (const pll::Partition::NniScore& m : moves)
            scored.push_back(IndexedScore{m.node->node_index, m.type, m.length, m.lnl});
          hood.set_value(std::move(scored));
        } catch (...) {
          hood.set_exception(std::current_exception());
          moves.clear();
        }
        if (!moves.empty()) WalkAhead(*task, *rep, moves);
      }
    } catch (...) {
      try {
        task->result.set_exception(std::current_exception());
      } catch (const std::future_error&) {   // already fulfilled: the failure came after the visit
      }
    }
    task.reset();
    std::lock_guard<std::mutex> lock(mutex_);
    pools_[rep->device()].free.push_back(std::move(rep));
    replica_back_.notify_one();
  }
}

} } // namespace pt::move_tester
