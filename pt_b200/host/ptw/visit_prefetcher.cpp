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

VisitPrefetcher::Batch::~Batch()
{
  if (parent) pll_utree_destroy(parent, pll::cb_erase_data);
  if (parent_first) pll_utree_destroy(parent_first, pll::cb_erase_data);
}

VisitPrefetcher::TreeHolder::~TreeHolder()
{
  if (t) pll_utree_destroy(t, pll::cb_erase_data);
}

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
  batches_.clear();
  batch_order_.clear();
  noted_.clear();
  pll::Partition::ClearVisits();
}

void VisitPrefetcher::Submit(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* node, int type,
                             const Authority& authority) noexcept
{
  if (!enabled()) return;
  authority_.store(&authority, std::memory_order_relaxed);
  try {
    NoteThreshold(authority);
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
  std::lock_guard<std::mutex> lock(mutex_);
  Noted& noted = noted_[&owner];
  if (noted.tree != tree) {   // moves of an earlier tree whose queue was never visited (nothing was granted)
    noted.tree = tree;
    noted.moves.clear();
  }
  noted.moves.emplace_back(node->node_index, type);
}

void VisitPrefetcher::Flush(pll::Partition& owner, pll_utree_t* tree, pll_unode_t* at, bool edge_invalid) noexcept
{
  if (!enabled()) return;
  std::vector<std::pair<unsigned int, int>> moves;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    auto it = noted_.find(&owner);
    if (it == noted_.end() || it->second.tree != tree) return;
    moves.swap(it->second.moves);
    it->second.tree = nullptr;
  }
  try {
    for (const std::pair<unsigned int, int>& m : moves)
      Enqueue(owner.device(), owner.GetModel(), tree, RecordByIndex(tree, m.first), m.second, 0, at, edge_invalid);
  } catch (const std::exception& e) {
    disabled_ = true;
    std::fprintf(stderr, "visit prefetcher: turned off (%s)\n", e.what());
  } catch (...) {
    disabled_ = true;
  }
}

bool VisitPrefetcher::Enqueue(int device, const pll::Model& model, pll_utree_t* tree, pll_unode_t* node, int type,
                              int depth, const pll_unode_t* at, bool edge_invalid)
{
  const unsigned int n_nodes = tree->tip_count + tree->inner_count;
  // Which state the visit starts from: the FIRST move of a queue is visited with the tree as the move
  // tester left it (oriented at the last node tested: `at`), every later one either after MoveBack's
  // FULL traversal at the virtual root (this task) or after a failed sibling (BuildSibling).
  std::shared_ptr<Batch> group;
  const std::pair<const pll_utree_t*, uint64_t> id(tree, pll::Partition::TreeFingerprint(tree));
  if (depth == 0) {
    std::lock_guard<std::mutex> lock(mutex_);
    std::shared_ptr<Batch>& slot = batches_[id];
    if (!slot) {
      slot = std::make_shared<Batch>();
      slot->id = ++batch_;
      batch_order_.push_back(id);
      while (batch_order_.size() > 4096) {
        batches_.erase(batch_order_.front());
        batch_order_.pop_front();
      }
    }
    group = batches_[id];
  }
  const bool first = group && group->moves.empty() && at != nullptr;
  const unsigned int orient_index = first ? at->node_index : GetVirtualRoot(tree)->node_index;
  const bool orient_invalid = first && edge_invalid;
  // a clone of `tree` with the flags of that state
  auto oriented_clone = [&](unsigned int index, bool invalid) -> pll_utree_t* {
    pll_utree_t* c = pll_utree_clone(tree);
    if (!c) throw std::runtime_error("pll_utree_clone failed");
    pll_utree_every(c, pll::cb_copy_clv_traversal);
    std::vector<pll_unode_t*> buffer(n_nodes, nullptr);
    unsigned int count = 0;
    pll_unode_t* r = RecordByIndex(c, index);
    if (!pll_utree_traverse(r, PLL_TREE_TRAVERSE_POSTORDER, pll::cb_full_traversal, buffer.data(), &count)) {
      pll_utree_destroy(c, pll::cb_erase_data);
      throw std::runtime_error("the tree's orientation node is not an inner node");
    }
    if (invalid) pll::InvalidateEdgeClvs(r);
    return c;
  };
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
    } parent{oriented_clone(orient_index, orient_invalid)};
    pll_unode_t* moved = RecordByIndex(parent.t, node->node_index);
    pll_utree_nni(moved, type, nullptr);
    pll::InvalidateEdgeClvs(moved);
    task->tree = pll_utree_clone(parent.t);
    if (!task->tree) throw std::runtime_error("pll_utree_clone failed");
    pll_utree_every(task->tree, pll::cb_copy_clv_traversal);
  }
  task->parent_root_index = orient_index;
  task->parent_edge_invalid = orient_invalid;
  const uint64_t content = pll::Partition::ExactContentId(id.second, orient_index);
  const uint64_t key = pll::Partition::VisitKeyFor(content, GetVirtualRoot(task->tree), n_nodes);
  if (!key) return false;
  task->device = device;
  task->model = model;
  task->depth = depth;
  task->node_index = node->node_index;
  task->type = type;
  // the group this move belongs to: the moves one QueueMoves call sends here, in order (depth 0 only)
  if (group) {
    const Authority* authority = authority_.load(std::memory_order_relaxed);
    std::lock_guard<std::mutex> lock(mutex_);
    if (!group->parent) {
      group->device = device;
      group->model = model;
      group->parent_root_index = GetVirtualRoot(tree)->node_index;
      // the parent tree with the flags a FULL traversal at its root leaves, and with the tester's
      group->parent = oriented_clone(group->parent_root_index, false);
      if (first) group->parent_first = oriented_clone(orient_index, orient_invalid);
    }
    // two moves of a queue can lead to the same topology (the root entry repeats its first child's
    // edge): the authority grants only the first (src/authority.cpp:139-152), so only that one counts
    if (authority && !group->topologies.insert(authority->GetKey(task->tree)).second) return false;
    task->group = group;
    task->index = task->run_start = group->moves.size();
    group->moves.emplace_back(node->node_index, type);
  }
  if (pll::Partition::VisitKnown(key)) return false;   // (a worker walking ahead got there first)
  pll::Partition::PublishVisit(key, task->result.get_future().share(),
                               sizeof(double) * (task->tree->tip_count + 3 * task->tree->inner_count));
  {
    std::lock_guard<std::mutex> lock(mutex_);
    task->batch = group ? group->id : ++batch_;
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
  if (!enabled()) return false;
  if (const Authority* authority = authority_.load(std::memory_order_relaxed)) NoteThreshold(*authority);
  if (!score_hoods_) return false;
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

void VisitPrefetcher::NoteThreshold(const Authority& authority)
{
  const double now = authority.GetThresholdScore();
  std::lock_guard<std::mutex> lock(mutex_);
  if (have_threshold_ && now == last_threshold_) return;
  const bool moved = have_threshold_;
  have_threshold_ = true;
  last_threshold_ = now;
  if (!moved || batches_.empty()) return;
  rescan_ = true;
  wake_.notify_one();
}

double VisitPrefetcher::RunVisit(Task& task, pll::Partition& rep, bool state_on_replica,
                                 const std::function<void(double)>& before_release)
{
  if (!SameModel(rep.GetModel(), task.model)) {
    rep.SetModel(task.model);
    state_on_replica = false;
  }
  if (!task.previous) {
    pll_unode_t* moved = RecordByIndex(task.tree, task.node_index);
    // back to the parent tree: its exact CLVs, oriented at its root, are what the wanderer's
    // partition holds when it starts this visit
    pll_utree_nni(moved, task.type, nullptr);
    pll_unode_t* oriented_at = RecordByIndex(task.tree, task.parent_root_index);
    rep.TraversalUpdate(oriented_at, pll::TraversalType::FULL);
    if (task.parent_edge_invalid) pll::InvalidateEdgeClvs(oriented_at);   // (the radius-0 tester's last state)
    // the move, as MoveForward applies it
    pll_utree_nni(moved, task.type, nullptr);
    pll::InvalidateEdgeClvs(moved);
  } else if (!state_on_replica ||
             rep.ContentId() != pll::Partition::ExactContentId(pll::Partition::TreeFingerprint(task.previous->t),
                                                               GetVirtualRoot(task.previous->t)->node_index)) {
    // the failed sibling's optimised tree, everything exact and oriented at its root: the state its
    // visit ended in (Partition::OptimizeAllBranches, TakePublishedVisit)
    rep.TraversalUpdate(GetVirtualRoot(task.previous->t), pll::TraversalType::FULL);
  }
  pll_unode_t* root = GetVirtualRoot(task.tree);
  rep.OptimizeAllBranches(root);
  pll::Partition::VisitResult res;
  const unsigned int n = task.tree->tip_count + task.tree->inner_count;
  res.lengths.assign(task.tree->tip_count + 3 * task.tree->inner_count, 0.0);
  for (unsigned int i = 0; i < n; ++i) {
    pll_unode_t* first = task.tree->nodes[i];
    pll_unode_t* r = first;
    do {
      if (r->node_index < res.lengths.size()) res.lengths[r->node_index] = r->length;
      r = r->next;
    } while (r && r != first);
  }
  // what the wanderer does next (src/wanderer.cpp:192-198); the CLVs already face the root
  rep.TraversalUpdate(root, pll::TraversalType::PARTIAL);
  const double lnl = rep.LogLikelihood(root);
  before_release(lnl);   // what must be announced before the wanderer can see this visit's result
  if (task.group) {
    const std::shared_ptr<Batch> group = task.group;
    const VisitId id(task.index, task.previous ? task.run_start : task.index);
    res.taken = [this, group, id, lnl] { Taken(group, id, lnl); };
  }
  task.result.set_value(std::move(res));
  return lnl;
}

// What the wanderer is about to decide about the tree it has just taken (src/wanderer.cpp:196-207,
// src/authority.cpp:191-199): final for this queue's path.  If the tree fails, the visit of the next
// move is registered before this returns -- the wanderer asks for it within microseconds.
void VisitPrefetcher::Taken(const std::shared_ptr<Batch>& group, VisitId id, double lnl)
{
  const Authority* authority = authority_.load(std::memory_order_relaxed);
  if (!authority || !enabled()) return;
  const bool good = lnl >= authority->GetThresholdScore();
  {
    std::lock_guard<std::mutex> lock(mutex_);
    if (stop_) return;
    group->verdict[id] = good;
  }
  if (good) return;
  try {
    StartJobs(group);
  } catch (...) {
  }
}

void VisitPrefetcher::StartJobs(const std::shared_ptr<Batch>& group)
{
  for (const Job& job : Scan(group)) {
    std::unique_ptr<Task> t = BuildSibling(job);
    if (!t) continue;
    std::lock_guard<std::mutex> lock(mutex_);
    queue_.push_back(std::move(t));
    wake_.notify_one();
  }
}

// The queue of `group` as the wanderer will walk it under the present threshold (see "Sibling
// chains" in the header): which visit follows which, as far as the finished visits tell.
std::vector<VisitPrefetcher::Job> VisitPrefetcher::Scan(const std::shared_ptr<Batch>& group)
{
  std::vector<Job> jobs;
  const Authority* authority = authority_.load(std::memory_order_relaxed);
  if (!group || !authority) return jobs;
  const double threshold = authority->GetThresholdScore();
  std::lock_guard<std::mutex> lock(mutex_);
  bool exact = true;
  size_t run_start = 0;
  for (size_t i = 0; i < group->moves.size(); ++i) {
    const auto it = group->done.find(exact ? VisitId(i, i) : VisitId(i, run_start));
    if (it == group->done.end()) break;   // not finished yet: its completion scans again
    const auto v = group->verdict.find(it->first);
    if (v != group->verdict.end() ? v->second : it->second.lnl >= threshold) {
      exact = true;   // good (so far): the wanderer descends, and MoveBack restores the exact parent
      continue;
    }
    if (exact) run_start = i;
    exact = false;
    if (i + 1 >= group->moves.size()) break;
    if (!group->started.insert(VisitId(i + 1, run_start)).second) continue;
    Job job;
    job.group = group;
    job.index = i + 1;
    job.run_start = run_start;
    job.previous = it->second.tree;
    job.run.assign(group->moves.begin() + run_start, group->moves.begin() + i + 1);
    job.move = group->moves[i + 1];
    jobs.push_back(std::move(job));
  }
  return jobs;
}

// the visit of job.move from the state the failed visit job.previous leaves behind: the parent
// tree's flags with the invalidated edge ends of every move of the present run of failures
std::unique_ptr<VisitPrefetcher::Task> VisitPrefetcher::BuildSibling(const Job& job)
{
  std::unique_ptr<Task> task(new Task);
  const pll_utree_t* failed = job.previous->t;
  const unsigned int n_nodes = failed->tip_count + failed->inner_count;
  {
    struct Holder {
      pll_utree_t* t;
      ~Holder() { if (t) pll_utree_destroy(t, pll::cb_erase_data); }
    // (a run of failures that starts with the queue's first visit still sees the tester's orientation;
    //  any later run starts after a good tree, i.e. after MoveBack's FULL traversal at the root)
    } parent{pll_utree_clone(job.run_start == 0 && job.group->parent_first ? job.group->parent_first
                                                                           : job.group->parent)};
    if (!parent.t) return nullptr;
    pll_utree_every(parent.t, pll::cb_copy_clv_traversal);
    // every failed sibling was applied, had its edge ends invalidated and was undone (src/wanderer.cpp:160-176)
    for (const std::pair<unsigned int, int>& m : job.run) pll::InvalidateEdgeClvs(RecordByIndex(parent.t, m.first));
    pll_unode_t* moved = RecordByIndex(parent.t, job.move.first);
    pll_utree_nni(moved, job.move.second, nullptr);
    pll::InvalidateEdgeClvs(moved);
    task->tree = pll_utree_clone(parent.t);
    if (!task->tree) return nullptr;
    pll_utree_every(task->tree, pll::cb_copy_clv_traversal);
  }
  const uint64_t content = pll::Partition::ExactContentId(pll::Partition::TreeFingerprint(failed),
                                                          GetVirtualRoot(job.previous->t)->node_index);
  const uint64_t key = pll::Partition::VisitKeyFor(content, GetVirtualRoot(task->tree), n_nodes);
  if (!key || pll::Partition::VisitKnown(key)) return nullptr;
  task->group = job.group;
  task->index = job.index;
  task->run_start = job.run_start;
  task->previous = job.previous;
  task->device = job.group->device;
  task->model = job.group->model;
  task->node_index = job.move.first;
  task->parent_root_index = job.group->parent_root_index;
  task->type = job.move.second;
  task->depth = 0;
  task->batch = job.group->id;
  pll::Partition::PublishVisit(key, task->result.get_future().share(),
                               sizeof(double) * (task->tree->tip_count + 3 * task->tree->inner_count));
  ++chained_;
  static const bool debug = std::getenv("PTGPU_PREFETCH_DEBUG") != nullptr;
  if (debug)
    std::fprintf(stderr, "[prefetch] chained key %016llx content %016llx batch %llu index %zu run_start %zu\n",
                 (unsigned long long)key, (unsigned long long)content, task->batch, task->index, task->run_start);
  return task;
}

void VisitPrefetcher::Work()
{
  static const bool debug = std::getenv("PTGPU_PREFETCH_DEBUG") != nullptr;
  for (;;) {
    std::unique_ptr<Task> task;
    std::unique_ptr<pll::Partition> rep;
    std::vector<std::shared_ptr<Batch>> rescan;
    {
      std::unique_lock<std::mutex> lock(mutex_);
      wake_.wait(lock, [this] { return stop_ || rescan_ || !queue_.empty(); });
      if (stop_) break;
      if (rescan_) {
        rescan_ = false;
        for (const auto& b : batches_) rescan.push_back(b.second);
      } else {
        // The wanderer walks depth first: the queue it fills last is the one it empties first
        // (move_queues_ is a stack, src/wanderer.cpp:122-131), and each queue from its front
        // (:157-158).  So: what the wanderer itself asked for before speculation, newest batch
        // first, in queue order within it.
        size_t pick = 0;
        for (size_t i = 1; i < queue_.size(); ++i) {
          const Task& a = *queue_[i];
          const Task& b = *queue_[pick];
          if (a.depth != b.depth ? a.depth < b.depth
                                 : (a.batch != b.batch ? a.batch > b.batch : a.index < b.index))
            pick = i;
        }
        task = std::move(queue_[pick]);
        queue_.erase(queue_.begin() + pick);
        Pool& pool = pools_[task->device];
        replica_back_.wait(lock, [&] { return stop_ || !pool.free.empty(); });
        if (stop_) break;   // the task's promise breaks: the wanderer computes the visit in place
        rep = std::move(pool.free.back());
        pool.free.pop_back();
      }
    }
    if (!rescan.empty()) {
      // the threshold has moved: visits found good before may have failed since
      std::sort(rescan.begin(), rescan.end(),
                [](const std::shared_ptr<Batch>& a, const std::shared_ptr<Batch>& b) { return a->id > b->id; });
      for (const std::shared_ptr<Batch>& group : rescan) {
        try {
          StartJobs(group);
        } catch (...) {
        }
      }
      continue;
    }
    // the visit; then, while the optimised tree is not a good one and the wanderer's path goes on
    // from it, the next move of the same queue on the same replica (whose content is that state)
    std::shared_ptr<TreeHolder> on_replica;   // the tree whose exact CLVs the replica holds
    while (task) {
      std::unique_ptr<Task> next;
      try {
        std::promise<Neighborhood> hood;
        bool good = true, score = false;
        const bool state_on_replica = task->previous && task->previous == on_replica;
        const double lnl = RunVisit(*task, *rep, state_on_replica, [&](double value) {
          const Authority* authority = authority_.load(std::memory_order_relaxed);
          good = !authority || value >= authority->GetThresholdScore();
          score = good && score_hoods_ && radius_ >= 0;
          if (!score) return;
          // the neighbourhood of the optimised tree, for the wanderer's QueueMoves: announced before
          // the visit's result is released, so the wanderer cannot get ahead of the announcement
          const uint64_t fp = pll::Partition::TreeFingerprint(task->tree);
          std::lock_guard<std::mutex> lock(mutex_);
          if (hoods_.emplace(fp, hood.get_future().share()).second) hood_order_.push_back(fp);
          while (hood_order_.size() > 65536) {
            hoods_.erase(hood_order_.front());
            hood_order_.pop_front();
          }
        });
        if (debug)
          std::fprintf(stderr, "[prefetch] visit %s batch %llu index %zu/%zu run_start %zu good %d\n",
                       !task->previous ? "A" : (state_on_replica ? "B" : "B'"), task->batch, task->index,
                       task->group ? task->group->moves.size() : 0, task->run_start, (int)good);
        on_replica.reset();
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
        if (task->group) {
          // keep the optimised tree: the next move of the queue may have to start from it
          std::shared_ptr<TreeHolder> kept = std::make_shared<TreeHolder>(task->tree);
          task->tree = nullptr;
          {
            std::lock_guard<std::mutex> lock(mutex_);
            task->group->done[VisitId(task->index, task->previous ? task->run_start : task->index)] = Done{lnl, kept};
          }
          // (the neighbourhood scoring works on directional buffers and leaves the per-node CLVs alone)
          on_replica = kept;
          for (const Job& job : Scan(task->group)) {
            std::unique_ptr<Task> t = BuildSibling(job);
            if (!t) continue;
            if (!next && t->previous == on_replica) {
              next = std::move(t);
              continue;
            }
            std::lock_guard<std::mutex> lock(mutex_);
            queue_.push_back(std::move(t));
            wake_.notify_one();
          }
        }
      } catch (...) {
        try {
          task->result.set_exception(std::current_exception());
        } catch (const std::future_error&) {   // already fulfilled: the failure came after the visit
        }
      }
      task = std::move(next);
      bool stopping;
      {
        std::lock_guard<std::mutex> lock(mutex_);
        stopping = stop_;
      }
      if (stopping) task.reset();
    }
    task.reset();
    std::lock_guard<std::mutex> lock(mutex_);
    pools_[rep->device()].free.push_back(std::move(rep));
    replica_back_.notify_one();
  }
}

} } // namespace pt::move_tester
