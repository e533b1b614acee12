// pt_b200/host/ptw/visit_prefetcher.cpp -- see the header.
#include "visit_prefetcher.hpp"

#include <cstdlib>
#include <exception>

#include <pll_util.hpp>

#include "ordered_tree.hpp"   // GetVirtualRoot

namespace pt { namespace move_tester {

VisitPrefetcher& VisitPrefetcher::Instance()
{
  static VisitPrefetcher instance;
  return instance;
}

VisitPrefetcher::VisitPrefetcher() {}

void VisitPrefetcher::Start(int workers)
{
  if (!workers_.empty()) return;
  int n = workers;
  if (const char* env = std::getenv("PTGPU_PREFETCH_VISITS")) n = std::atoi(env);
  {
    std::lock_guard<std::mutex> lock(mutex_);
    stop_ = false;
  }
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
  for (std::thread& t : workers_)
    if (t.joinable()) t.join();
  workers_.clear();
  free_replicas_.clear();   // the engines go here, while the CUDA runtime is still up
  queue_.clear();
  pll::Partition::ClearVisits();
}

void VisitPrefetcher::Submit(pll::Partition& owner, pll_utree_t* tree)
{
  if (!enabled()) return;
  const uint64_t fp = pll::Partition::TreeFingerprint(tree);
  if (pll::Partition::VisitKnown(fp)) return;
  std::unique_ptr<Task> task(new Task);
  task->owner = &owner;
  // what Wanderer::MoveForward does before it optimises (src/wanderer.cpp:163-165)
  task->tree = pll_utree_clone(tree);
  pll_utree_every(task->tree, pll::cb_copy_clv_traversal);
  pll::Partition::PublishVisit(fp, task->result.get_future().share());
  bool known_owner;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    known_owner = free_replicas_.count(&owner) != 0;
  }
  if (!known_owner) {
    std::vector<std::unique_ptr<pll::Partition>> made;
    for (size_t i = 0; i < workers_.size(); ++i) made.push_back(owner.Replica());
    std::lock_guard<std::mutex> lock(mutex_);
    free_replicas_[&owner] = std::move(made);
  }
  {
    std::lock_guard<std::mutex> lock(mutex_);
    const pll_utree_t*& last = last_tree_[&owner];
    if (last != tree) { last = tree; ++batch_; }
    task->batch = batch_;
    queue_.push_back(std::move(task));
    ++submitted_;
  }
  wake_.notify_one();
}

void VisitPrefetcher::Work()
{
  for (;;) {
    std::unique_ptr<Task> task;
    std::unique_ptr<pll::Partition> rep;
    {
      std::unique_lock<std::mutex> lock(mutex_);
      wake_.wait(lock, [this] { return stop_ || !queue_.empty(); });
      if (queue_.empty()) break;   // stop_ and nothing left
      // The wanderer walks depth first: the queue it fills last is the one it empties first
      // (move_queues_ is a stack, src/wanderer.cpp:122-131), and each queue from its front
      // (:157-158).  So: newest batch first, oldest submission within it.
      size_t pick = 0;
      for (size_t i = 1; i < queue_.size(); ++i)
        if (queue_[i]->batch > queue_[pick]->batch) pick = i;
      task = std::move(queue_[pick]);
      queue_.erase(queue_.begin() + pick);
      // as many replicas per owner as workers: one is always free
      std::vector<std::unique_ptr<pll::Partition>>& pool = free_replicas_[task->owner];
      rep = std::move(pool.back());
      pool.pop_back();
    }
    try {
      pll_unode_t* root = GetVirtualRoot(task->tree);
      // the replica's CLV buffers hold some other tree: start from a FULL traversal; the result
      // of the optimisation does not depend on which CLVs happened to be valid at the start
      rep->TraversalUpdate(root, pll::TraversalType::FULL);
      rep->OptimizeAllBranches(root);
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
      pll_utree_destroy(task->tree, pll::cb_erase_data);
      task->tree = nullptr;
      task->result.set_value(std::move(res));
    } catch (...) {
      if (task->tree) pll_utree_destroy(task->tree, pll::cb_erase_data);
      task->result.set_exception(std::current_exception());
    }
    std::lock_guard<std::mutex> lock(mutex_);
    free_replicas_[task->owner].push_back(std::move(rep));
  }
}

} } // namespace pt::move_tester
