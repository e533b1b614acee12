// pt_b200/host/src/pll_partition.cpp -- pt::pll::Partition over the B200 engine.
//
// Host control flow follows the reference's lib/libptpll/src/pll_partition.cpp
// (cited as PTPLL/pll_partition.cpp:line) so that callers observe the same
// sequence of states; every numeric step is an extern "C" call into
// libptgpu.so (include/ptgpu.h).  Nothing here computes a likelihood on the
// CPU: if the engine cannot be created the constructor throws.
#include "pll_partition.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <thread>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <functional>
#include <future>
#include <map>
#include <mutex>
#include <unordered_map>
#include <stdexcept>
#include <string>
#include <strings.h>
#include <vector>

#include "ptgpu.h"

static_assert(sizeof(pll_operation_t) == sizeof(ptgpu_operation_t),
              "ptgpu_operation_t must mirror pll_operation_t");

namespace pt { namespace pll {

namespace {

// pll-modules' DNA model names (PLLMOD/util/models_dna.c:64-106); SetModel
// accepts exactly these, case-insensitively, and GetModel reports the
// canonical spelling, as pllmod_util_model_info_dna does.
const char* const kDnaModels[] = {
    "JC", "K80", "F81", "HKY", "TN93ef", "TN93", "K81", "K81uf", "TPM2", "TPM2uf", "TPM3",
    "TPM3uf", "TIM1", "TIM1uf", "TIM2", "TIM2uf", "TIM3", "TIM3uf", "TVMef", "TVM", "SYM", "GTR"};

const char* canonical_model_name(const std::string& name)
{
  for (const char* m : kDnaModels)
    if (strcasecmp(name.c_str(), m) == 0) return m;
  return nullptr;
}

const double kMapLambda = 10.0;  // exponential prior rate, PTPLL/pll_partition.cpp:255,318,435

std::atomic<unsigned int> g_next_device{0};

int pick_device()
{
  if (const char* env = std::getenv("PTGPU_DEVICE")) return std::atoi(env);
  const int n = ptgpu_device_count();
  if (n <= 0) return 0;  // ptgpu_create reports the missing device
  return static_cast<int>(g_next_device.fetch_add(1) % static_cast<unsigned int>(n));
}

int visit_all(pll_unode_t*) { return 1; }

// PTGPU_HOST_STATS=1: wall time and call counts of the Partition entry points,
// printed when a Partition dies (where does a ptw run spend its time?)
struct HostStat { std::atomic<unsigned long long> ns{0}, calls{0}; };
enum { kStatTraversal, kStatOptimizeBranch, kStatLogLikelihood, kStatNeighborhood, kStatAllBranches, kStatVisitWait, kStatCount };
HostStat g_stats[kStatCount];
const char* const kStatNames[kStatCount] = {"TraversalUpdate", "OptimizeBranch", "LogLikelihood",
                                            "ScoreNniNeighborhood", "OptimizeAllBranches(total)",
                                            "  waiting for a prefetched visit"};
const bool g_stats_on = std::getenv("PTGPU_HOST_STATS") != nullptr;

class StatScope {
 public:
  // `counted` = false for the background replicas: the statistics are the wanderers' own time
  explicit StatScope(int which, bool counted = true) : which_(which), on_(g_stats_on && counted)
  {
    if (on_) t0_ = std::chrono::steady_clock::now();
  }
  ~StatScope()
  {
    if (!on_) return;
    const auto dt = std::chrono::steady_clock::now() - t0_;
    g_stats[which_].ns += std::chrono::duration_cast<std::chrono::nanoseconds>(dt).count();
    g_stats[which_].calls += 1;
  }
 private:
  int which_;
  bool on_;
  std::chrono::steady_clock::time_point t0_;
};

void PrintStats()
{
  if (!g_stats_on) return;
  for (int k = 0; k < kStatCount; ++k)
    std::fprintf(stderr, "[ptgpu host] %-28s %10llu calls %10.3f s\n", kStatNames[k],
                 g_stats[k].calls.load(), g_stats[k].ns.load() * 1e-9);
}

} // namespace

void Partition::Check(int rc) const
{
  if (rc == PTGPU_OK) return;
  const std::string msg = ptgpu_last_error();
  if (rc == PTGPU_ERR_NOT_FINITE) throw std::range_error(msg);
  if (rc == PTGPU_ERR_ARG || rc == PTGPU_ERR_STATE) throw std::invalid_argument(msg);
  throw std::runtime_error("ptgpu: " + msg);
}

// ---- site shards ---------------------------------------------------------------------------
// One worker thread per extra shard; Run(fn) executes fn(0) on the caller's thread and fn(s) on
// worker s-1, and returns when all are done.  A shard's kernel waits for its peers' kernels, so
// the calls must be in flight together; the engine library keeps its error text per thread,
// hence every thread reports its own.
class Partition::ShardPool {
 public:
  explicit ShardPool(unsigned int shards) : rc_(shards, 0), msg_(shards)
  {
    for (unsigned int s = 1; s < shards; ++s) workers_.emplace_back([this, s] { Loop(s); });
  }
  ~ShardPool()
  {
    {
      std::lock_guard<std::mutex> lock(m_);
      quit_ = true;
    }
    wake_.notify_all();
    for (std::thread& t : workers_) t.join();
  }
  // returns the first failing shard's (code, message), or (0, "")
  std::pair<int, std::string> Run(const std::function<int(unsigned int)>& fn)
  {
    {
      std::lock_guard<std::mutex> lock(m_);
      fn_ = &fn;
      pending_ = static_cast<unsigned int>(workers_.size());
      ++generation_;
    }
    wake_.notify_all();
    Invoke(fn, 0);
    {
      std::unique_lock<std::mutex> lock(m_);
      done_.wait(lock, [this] { return pending_ == 0; });
      fn_ = nullptr;
    }
    for (size_t s = 0; s < rc_.size(); ++s)
      if (rc_[s] != PTGPU_OK) return std::make_pair(rc_[s], msg_[s]);
    return std::make_pair(PTGPU_OK, std::string());
  }

 private:
  void Invoke(const std::function<int(unsigned int)>& fn, unsigned int s)
  {
    try {
      rc_[s] = fn(s);
      msg_[s] = rc_[s] == PTGPU_OK ? std::string() : std::string(ptgpu_last_error());
    } catch (const std::exception& ex) {
      rc_[s] = PTGPU_ERR_STATE;
      msg_[s] = ex.what();
    }
  }
  void Loop(unsigned int s)
  {
    unsigned long long seen = 0;
    for (;;) {
      const std::function<int(unsigned int)>* fn = nullptr;
      {
        std::unique_lock<std::mutex> lock(m_);
        wake_.wait(lock, [&] { return quit_ || generation_ != seen; });
        if (quit_) return;
        seen = generation_;
        fn = fn_;
      }
      Invoke(*fn, s);
      {
        std::lock_guard<std::mutex> lock(m_);
        --pending_;
      }
      done_.notify_one();
    }
  }
  std::mutex m_;
  std::condition_variable wake_, done_;
  const std::function<int(unsigned int)>* fn_ = nullptr;
  unsigned long long generation_ = 0;
  unsigned int pending_ = 0;
  bool quit_ = false;
  std::vector<int> rc_;
  std::vector<std::string> msg_;
  std::vector<std::thread> workers_;
};

namespace {
// Sharded operations of different Partitions are not interleaved: each is a set of kernels
// that spin on one another, and two such sets that each hold part of a device while waiting
// for the rest of it would wait for ever.  (A sharded alignment is one that fills the GPUs.)
std::mutex g_sharded_call;

void ThrowFor(int rc, const std::string& msg)
{
  if (rc == PTGPU_ERR_NOT_FINITE) throw std::range_error(msg);
  if (rc == PTGPU_ERR_ARG || rc == PTGPU_ERR_STATE) throw std::invalid_argument(msg);
  throw std::runtime_error("ptgpu: " + msg);
}
} // namespace

void Partition::OnShards(const std::function<int(ptgpu_engine*, unsigned int)>& call) const
{
  if (shards_.size() <= 1) {
    Check(call(engine_, 0));
    return;
  }
  std::lock_guard<std::mutex> lock(g_sharded_call);
  const auto failed = pool_->Run([&](unsigned int s) { return call(shards_[s], s); });
  if (failed.first != PTGPU_OK) ThrowFor(failed.first, failed.second);
}

// the engines: one, or `wanted_shards` with contiguous tile-aligned site ranges (fewer if the
// alignment has fewer tiles), on consecutive devices starting at device_
void Partition::CreateEngines(unsigned int clv_buffers, unsigned int scale_buffers,
                              unsigned int rate_cats, unsigned int wanted_shards)
{
  const unsigned int kTile = 256;  // the engine's reduction tile (kernels.cuh kTileSites)
  const unsigned int tiles = (site_count_ + kTile - 1) / kTile;
  unsigned int n = std::max(1u, std::min(std::min(wanted_shards, 8u), tiles));
  const unsigned int per = (tiles + n - 1) / n;
  n = (tiles + per - 1) / per;   // no empty shard
  shard_begin_.assign(1, 0u);
  for (unsigned int s = 0; s < n; ++s)
    shard_begin_.push_back(std::min<unsigned long long>(site_count_, 1ull * (s + 1) * per * kTile));
  const int ndev = std::max(1, ptgpu_device_count());
  try {
    for (unsigned int s = 0; s < n; ++s) {
      ptgpu_engine* e = nullptr;
      Check(ptgpu_create(tip_node_count_, clv_buffers, shard_begin_[s + 1] - shard_begin_[s], rate_cats,
                         branch_count(), scale_buffers, (device_ + static_cast<int>(s)) % ndev, &e));
      shards_.push_back(e);
      if (n > 1) Check(ptgpu_set_shard(e, shard_begin_[s], site_count_));
    }
    if (n > 1) {
      Check(ptgpu_comm_init(shards_.data(), n));
      pool_ = std::make_shared<ShardPool>(n);
    }
  } catch (...) {
    DestroyEngines();
    throw;
  }
  engine_ = shards_[0];
}

void Partition::DestroyEngines()
{
  pool_.reset();
  // no shard may still be writing into a peer's mailbox when that peer goes away
  for (ptgpu_engine* e : shards_) ptgpu_synchronize(e);
  for (ptgpu_engine* e : shards_) ptgpu_destroy(e);
  shards_.clear();
  engine_ = nullptr;
}

// PTPLL/pll_partition.cpp:25-65
Partition::Partition(pll_utree_t* tree,
                     const Model& model,
                     const std::vector<std::string>& labels,
                     const std::vector<std::string>& sequences,
                     bool map_mode) :
    tip_node_count_(tree->tip_count),
    map_mode_(map_mode),
    engine_(nullptr),
    site_count_(0),
    device_(0),
    directional_base_(0)
{
  if (!TreeHealthy(tree)) throw std::invalid_argument("Tree is missing branch lengths");
  if (sequences.empty()) throw std::invalid_argument("No sequences supplied");
  site_count_ = static_cast<unsigned int>(sequences[0].size());

  // One CLV + scaler buffer per inner node as in pll_partition_create
  // (PTPLL/pll_partition.cpp:47-57), plus one per inner-node RECORD (3(n-2))
  // for the directional CLVs of the batched NNI scorer.
  const bool directional = !std::getenv("PTGPU_NO_DIRECTIONAL");
  const unsigned int extra = directional ? 3 * inner_node_count() : 0;
  directional_base_ = directional ? inner_node_count() : 0;
  // OptimizeAllBranches is ONE persistent kernel (ptgpu_optimize_all_branches) at every size: 4 / 54 /
  // 252 ms per call at 2k / 100k / 1M sites x 64 taxa against 100 / 113 / 364 ms for the reference's
  // call-by-call sequence (which round 1 preferred above 500k sites, before the reduction tree gave a
  // single branch's sweep enough CTAs).  PTGPU_STEPWISE_ALL_BRANCHES=1 or a site limit
  // (PTGPU_FUSED_ALL_BRANCHES_MAX_SITES) bring the call-by-call form back for A/B runs.
  {
    unsigned long limit = ~0ul;
    if (const char* env = std::getenv("PTGPU_FUSED_ALL_BRANCHES_MAX_SITES")) limit = std::strtoul(env, nullptr, 10);
    stepwise_all_branches_ = std::getenv("PTGPU_STEPWISE_ALL_BRANCHES") != nullptr || site_count_ > limit;
  }
  // PTGPU_COMPRESS_PATTERNS=1 (off by default, as pt never compresses): the engine
  // holds the alignment's unique columns and weights every sum over sites by their
  // multiplicities (pll_compress_site_patterns + pll_set_pattern_weights, LIBPLL/compress.c:138,
  // LIBPLL/pll.c:1047).  Scores then differ from the uncompressed ones in the last bits only
  // (another summation order; well inside the 1e-9 the tables are compared at).
  std::vector<std::string> unique_columns;
  std::vector<unsigned int> weights;
  const std::vector<std::string>* tip_sequences = &sequences;
  if (std::getenv("PTGPU_COMPRESS_PATTERNS") && site_count_ > 0) {
    bool ragged = false;   // left to SetTipStates, which names the offending sequence
    for (const std::string& s : sequences) ragged = ragged || s.size() != site_count_;
    if (!ragged) {
      unique_columns = sequences;
      std::vector<char*> rows;
      for (std::string& s : unique_columns) rows.push_back(&s[0]);
      unsigned int length = site_count_;
      weights.resize(length);
      Check(ptgpu_compress_site_patterns(rows.data(), static_cast<unsigned int>(rows.size()), &length,
                                         weights.data()));
      for (std::string& s : unique_columns) s.resize(length);
      weights.resize(length);
      site_count_ = length;
      tip_sequences = &unique_columns;
    }
  }
  // (PTGPU_HOST_STATS: where a process's start-up goes)
  static const bool show = std::getenv("PTGPU_HOST_STATS") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto mark = [&](const char* what) {
    const auto now = std::chrono::steady_clock::now();
    if (show)
      std::fprintf(stderr, "[ptgpu host] Partition(): %-24s %8.3f s\n", what,
                   std::chrono::duration<double>(now - t_last).count());
    t_last = now;
  };
  device_ = pick_device();
  mark("device (CUDA start-up)");
  unsigned int wanted_shards = 1;
  if (const char* env = std::getenv("PTGPU_SITE_SHARDS")) wanted_shards = std::max(1, std::atoi(env));
  CreateEngines(inner_node_count() + extra, inner_node_count() + extra,
                static_cast<unsigned int>(model.category_rates.size()), wanted_shards);
  mark("engine");
  // the whole-tree step program on a shard is ONE persistent kernel per GPU; the call-by-call
  // form would pay a launch and an exchange per Newton step
  if (shards_.size() > 1 && !std::getenv("PTGPU_STEPWISE_ALL_BRANCHES")) stepwise_all_branches_ = false;
  try {
    SetModel(model);
    mark("model");
    SetTipStates(tree, labels, *tip_sequences);
    mark("tips");
    if (!weights.empty())
      OnShards([&](ptgpu_engine* e, unsigned int s) {
        return ptgpu_set_pattern_weights(e, weights.data() + shard_begin_[s]);
      });
  } catch (...) {
    DestroyEngines();
    throw;
  }
  AllocateScratchBuffers();
}

// ---- visits computed ahead of time (see the header) -------------------------------------
namespace {

std::mutex g_visit_mutex;
std::unordered_map<uint64_t, std::shared_future<Partition::VisitResult>> g_visits;
std::deque<std::pair<uint64_t, size_t>> g_visit_order;   // publication order with each entry's bytes
size_t g_visit_bytes = 0;                                 // sum over g_visit_order (taken entries leave it late)
std::atomic<bool> g_visits_used{false};
std::atomic<unsigned long long> g_visits_served{0}, g_visits_computed{0};

inline uint64_t record_hash(const pll_unode_t* r)
{
  uint64_t bits;
  std::memcpy(&bits, &r->length, sizeof bits);
  uint64_t h = 0x9e3779b97f4a7c15ull * (r->node_index + 1);
  h ^= 0xc2b2ae3d27d4eb4full * (r->back->node_index + 1) + (h << 6) + (h >> 2);
  h ^= bits + 0x165667b19e3779f9ull + (h << 6) + (h >> 2);
  h ^= h >> 29; h *= 0xbf58476d1ce4e5b9ull; h ^= h >> 32;
  return h;
}

inline uint64_t node_hash(const pll_unode_t* n)
{
  uint64_t h = 0;
  const pll_unode_t* r = n;
  do {
    h += record_hash(r);   // commutative: the order in which records are met does not matter
    r = r->next;
  } while (r && r != n);
  return h;
}

} // namespace

uint64_t Partition::TreeFingerprint(const pll_utree_t* tree)
{
  uint64_t h = 0;
  for (unsigned int i = 0; i < tree->tip_count + tree->inner_count; ++i) h += node_hash(tree->nodes[i]);
  return h;
}

uint64_t Partition::FingerprintFrom(pll_unode_t* root)
{
  unsigned int found = 0;
  if (!pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, visit_all, travbuffer_.data(), &found))
    return 0;
  if (found != node_count()) return 0;
  uint64_t h = 0;
  for (unsigned int i = 0; i < found; ++i) h += node_hash(travbuffer_[i]);
  return h;
}

uint64_t Partition::ExactContentId(uint64_t tree_fingerprint, unsigned int root_index)
{
  uint64_t h = tree_fingerprint ^ (0x9e3779b97f4a7c15ull * (root_index + 1));
  h ^= h >> 31; h *= 0x7fb5d329728ea185ull; h ^= h >> 27; h *= 0x81dadef4bc2dd44dull; h ^= h >> 33;
  return h ? h : 1;
}

// start state of OptimizeAllBranches(root): content, tree (topology + lengths), the per-record
// orientation flags (a record without flags yet behaves like an invalid one), the root
uint64_t Partition::VisitKeyFor(uint64_t content_id, pll_unode_t* root, unsigned int node_count)
{
  if (!content_id) return 0;
  std::vector<pll_unode_t*> nodes(node_count, nullptr);
  unsigned int found = 0;
  if (!pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, visit_all, nodes.data(), &found)) return 0;
  if (found != node_count) return 0;
  uint64_t tree = 0, flags = 0;
  for (unsigned int i = 0; i < found; ++i) {
    tree += node_hash(nodes[i]);
    pll_unode_t* r = nodes[i];
    if (!r->next) continue;
    do {
      const int valid = r->data ? static_cast<node_info_t*>(r->data)->clv_valid : 0;
      if (valid) {
        uint64_t h = 0xd6e8feb86659fd93ull * (r->node_index + 1);
        h ^= h >> 32; h *= 0xd6e8feb86659fd93ull; h ^= h >> 32;
        flags += h;
      }
      r = r->next;
    } while (r != nodes[i]);
  }
  uint64_t h = content_id;
  h ^= tree + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
  h ^= flags + 0xc2b2ae3d27d4eb4full + (h << 6) + (h >> 2);
  h ^= 0x165667b19e3779f9ull * (root->node_index + 1) + (h << 6) + (h >> 2);
  h ^= h >> 29; h *= 0xbf58476d1ce4e5b9ull; h ^= h >> 32;
  static const bool debug = std::getenv("PTGPU_PREFETCH_DEBUG") != nullptr;
  if (debug)
    std::fprintf(stderr, "[prefetch] key %016llx = content %016llx tree %016llx flags %016llx root %u\n",
                 (unsigned long long)(h ? h : 1), (unsigned long long)content_id, (unsigned long long)tree,
                 (unsigned long long)flags, root->node_index);
  return h ? h : 1;
}

uint64_t Partition::VisitKey(pll_unode_t* root) const
{
  return VisitKeyFor(content_id_, root, node_count());
}

void Partition::PublishVisit(uint64_t key, std::shared_future<VisitResult> result, size_t bytes)
{
  std::lock_guard<std::mutex> lock(g_visit_mutex);
  // Results nobody came for (a predicted start state that did not occur) are dropped oldest first
  // once the registry holds more than its budget: a depth-first walk returns to the later moves of
  // an early queue only after everything below it, so the budget is large (256 MB of lengths;
  // PTGPU_VISIT_REGISTRY_MB) -- results dropped too early are simply computed in place again.
  std::deque<std::pair<uint64_t, size_t>>& order = g_visit_order;
  size_t& held = g_visit_bytes;
  static const size_t budget = [] {
    const char* env = std::getenv("PTGPU_VISIT_REGISTRY_MB");
    return (env ? std::strtoull(env, nullptr, 10) : 256ull) << 20;
  }();
  bytes += 256;   // the entry itself
  if (g_visits.emplace(key, std::move(result)).second) {
    order.emplace_back(key, bytes);
    held += bytes;
  }
  while (order.size() > 1 && (held > budget || order.size() > (1u << 20))) {
    g_visits.erase(order.front().first);
    held -= order.front().second;
    order.pop_front();
  }
  g_visits_used = true;
}

bool Partition::VisitKnown(uint64_t key)
{
  std::lock_guard<std::mutex> lock(g_visit_mutex);
  return g_visits.count(key) != 0;
}

void Partition::ClearVisits()
{
  std::lock_guard<std::mutex> lock(g_visit_mutex);
  g_visits.clear();
  g_visit_order.clear();
  g_visit_bytes = 0;
}

std::pair<unsigned long long, unsigned long long> Partition::VisitStatistics()
{
  return std::make_pair(g_visits_served.load(), g_visits_computed.load());
}

size_t Partition::ReplicaBytes(bool directional) const
{
  const size_t cats = model_.category_rates.size();
  // per-node (+ directional) CLVs and scalers, sumtable, the step-program kernel's scratch of the
  // same order; with directional buffers also one sumtable per NNI candidate
  const size_t clvs = static_cast<size_t>(inner_node_count()) * (directional ? 4 : 1) + 3 +
                      (directional ? 2 * inner_node_count() : 0);
  return clvs * site_count_ * (cats * 32 + 4) + (64u << 20);
}

// the registered result of OptimizeAllBranches for exactly this start state, if there is one:
// lengths into the tree, then the CLVs and matrices of the optimised tree by one FULL traversal
// (the state the computation itself ends in: everything exact, oriented at `root`)
bool Partition::TakePublishedVisit(uint64_t key, pll_unode_t* root)
{
  std::shared_future<VisitResult> pending;
  {
    std::lock_guard<std::mutex> lock(g_visit_mutex);
    auto it = g_visits.find(key);
    if (it == g_visits.end()) return false;
    pending = it->second;
    g_visits.erase(it);
  }
  VisitResult res;
  try {
    StatScope wait(kStatVisitWait, !is_replica_);
    res = pending.get();
  } catch (...) {
    return false;   // the background visit failed: compute here, where the caller sees the error
  }
  unsigned int found = 0;
  if (!pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, visit_all, travbuffer_.data(), &found) ||
      found != node_count())
    return false;
  for (unsigned int i = 0; i < found; ++i) {
    pll_unode_t* n = travbuffer_[i];
    pll_unode_t* r = n;
    do {
      if (r->node_index >= res.lengths.size()) return false;
      r = r->next;
    } while (r && r != n);
  }
  for (unsigned int i = 0; i < found; ++i) {
    pll_unode_t* n = travbuffer_[i];
    pll_unode_t* r = n;
    do {
      r->length = res.lengths[r->node_index];
      r = r->next;
    } while (r && r != n);
  }
  TraversalUpdate(root, TraversalType::FULL);
  if (res.taken) res.taken();
  return true;
}

Partition::Partition(const Partition& proto, ReplicaTag, bool directional) :
    tip_node_count_(proto.tip_node_count_),
    map_mode_(proto.map_mode_),
    engine_(nullptr),
    site_count_(proto.site_count_),
    device_(proto.device_),
    directional_base_(0),
    stepwise_all_branches_(proto.stepwise_all_branches_),
    is_replica_(true)
{
  if (proto.shards_.size() > 1)
    throw std::logic_error("Replica() of a site-sharded partition is not supported");
  const unsigned int extra = directional ? 3 * inner_node_count() : 0;
  directional_base_ = directional ? inner_node_count() : 0;
  CreateEngines(inner_node_count() + extra, inner_node_count() + extra,
                static_cast<unsigned int>(proto.model_.category_rates.size()), 1);
  try {
    SetModel(proto.model_);
    Check(ptgpu_copy_tip_states(engine_, proto.engine_));
  } catch (...) {
    DestroyEngines();
    throw;
  }
  AllocateScratchBuffers();
}

std::unique_ptr<Partition> Partition::Replica(bool directional) const
{
  return std::unique_ptr<Partition>(new Partition(*this, ReplicaTag{}, directional));
}

// PTPLL/pll_partition.cpp:67-80: the engine moves, scratch is re-allocated
Partition::Partition(Partition&& rhs) :
    tip_node_count_(rhs.tip_node_count_),
    map_mode_(rhs.map_mode_),
    engine_(rhs.engine_),
    model_(std::move(rhs.model_)),
    site_count_(rhs.site_count_),
    device_(rhs.device_),
    directional_base_(rhs.directional_base_),
    stepwise_all_branches_(rhs.stepwise_all_branches_),
    is_replica_(rhs.is_replica_),
    shards_(std::move(rhs.shards_)),
    shard_begin_(std::move(rhs.shard_begin_)),
    pool_(std::move(rhs.pool_))
{
  rhs.engine_ = nullptr;
  rhs.shards_.clear();
  AllocateScratchBuffers();
}

Partition::~Partition()
{
  if (engine_) {
    PrintStats();
    DestroyEngines();
  }
}

void Partition::AllocateScratchBuffers()
{
  travbuffer_.resize(node_count());
  branch_lengths_.resize(branch_count());
  matrix_indices_.resize(branch_count());
  operations_.resize(inner_node_count());
}

// PTPLL/pll_partition.cpp:87-108
void Partition::SetModel(const Model& model)
{
  const char* name = canonical_model_name(model.model_name);
  if (!name) throw std::invalid_argument("Invalid model name " + model.model_name);
  if (model.frequencies.size() != STATES || model.subst_params.size() != STATES * (STATES - 1) / 2)
    throw std::invalid_argument("Invalid number of frequencies or substitution rates");
  // (the reference sets frequencies and rates before it checks the category
  // count; the engine takes all three at once, so the check comes first)
  if (model_.category_rates.size() && model.category_rates.size() != model_.category_rates.size())
    throw std::invalid_argument("Invalid number of category rates");
  content_id_ = 0;
  OnShards([&](ptgpu_engine* e, unsigned int) {
    return ptgpu_set_model(e, model.frequencies.data(), model.subst_params.data(),
                           model.category_rates.data());
  });
  model_ = model;
  model_.model_name = name;
}

// PTPLL/pll_partition.cpp:110-151
Model Partition::GetModel() const { return model_; }

// PTPLL/pll_partition.cpp:153-194
void Partition::SetTipStates(pll_utree_t* tree,
                             const std::vector<std::string>& labels,
                             const std::vector<std::string>& sequences)
{
  if (labels.size() != sequences.size())
    throw std::invalid_argument("Number of labels does not match number of sequences");
  if (tip_node_count_ != labels.size())
    throw std::invalid_argument("Unexpected number of tip nodes supplied");

  std::map<std::string, unsigned int> clv_of_label;
  for (unsigned int i = 0; i < tip_node_count_; ++i) {
    const pll_unode_t* tip = tree->nodes[i];
    if (!clv_of_label.emplace(tip->label, tip->clv_index).second)
      throw std::invalid_argument(std::string("Error inserting tip label ") + tip->label +
                                  " into map (possibly a duplicate)");
  }
  std::vector<unsigned int> tips(tip_node_count_);
  std::vector<const char*> raw(tip_node_count_);
  for (unsigned int i = 0; i < tip_node_count_; ++i) {
    auto it = clv_of_label.find(labels[i]);
    if (it == clv_of_label.end())
      throw std::invalid_argument("Sequence with header " + labels[i] +
                                  " does not appear in the tree");
    if (sequences[i].size() != site_count_)
      throw std::invalid_argument("Sequence " + labels[i] + " has a different length");
    tips[i] = it->second;
    raw[i] = sequences[i].c_str();
  }
  // the reference's per-tip pll_set_tip_states loop, as one upload + one encoding kernel
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    std::vector<const char*> slice(raw);
    for (const char*& p : slice) p += shard_begin_[s];   // this shard's columns of every row
    return ptgpu_set_tip_states_batch(e, tip_node_count_, tips.data(), slice.data());
  });
}

// sum over all branches of log(lambda) - lambda*length in post-order, the edge
// of `tree` counted once (PTPLL/pll_partition.cpp:317-343)
double Partition::MapLogPrior(pll_unode_t* tree)
{
  std::vector<pll_unode_t*> nodes(node_count(), nullptr);
  unsigned int found = 0;
  if (!pll_utree_traverse(tree, PLL_TREE_TRAVERSE_POSTORDER, visit_all, nodes.data(), &found))
    throw std::invalid_argument("LogLikelihood() requires an inner node");
  if (found != nodes.size()) throw std::invalid_argument("Unexpected number of nodes");
  double prior = 0.0;
  for (pll_unode_t* n : nodes) {
    if (n == tree->back) continue;
    prior += std::log(kMapLambda) - (kMapLambda * n->length);
  }
  return prior;
}

// PTPLL/pll_partition.cpp:309-346
double Partition::LogLikelihood(pll_unode_t* tree, double* per_site_lnl)
{
  StatScope stat(kStatLogLikelihood, !is_replica_);
  double lnl = 0.0;
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    double mine = 0.0;   // the same global sum on every shard
    const int rc = ptgpu_compute_edge_loglikelihood(
        e, tree->clv_index, tree->scaler_index, tree->back->clv_index, tree->back->scaler_index,
        tree->pmatrix_index, per_site_lnl ? per_site_lnl + shard_begin_[s] : nullptr, &mine);
    if (s == 0) lnl = mine;
    return rc;
  });
  if (map_mode_) {
    // the reference adds the terms one by one to lnl, in this order
    std::vector<pll_unode_t*> nodes(node_count(), nullptr);
    unsigned int found = 0;
    if (!pll_utree_traverse(tree, PLL_TREE_TRAVERSE_POSTORDER, visit_all, nodes.data(), &found))
      throw std::invalid_argument("LogLikelihood() requires an inner node");
    if (found != nodes.size()) throw std::invalid_argument("Unexpected number of nodes");
    for (pll_unode_t* n : nodes) {
      if (n == tree->back) continue;
      lnl += std::log(kMapLambda) - (kMapLambda * n->length);
    }
  }
  return lnl;
}

// The Laplace approximation lives in lesplace (GSL), outside the hot path
// (--marginal is off in every BASELINE config).
double Partition::LogMarginalLikelihood(pll_unode_t*)
{
  throw std::logic_error("LogMarginalLikelihood is not provided by the B200 engine build "
                         "(lesplace/GSL path, out of scope)");
}

// PTPLL/pll_partition.cpp:348-393
unsigned int Partition::TraversalUpdate(pll_unode_t* root, TraversalType type)
{
  if (type == TraversalType::FULL) pending_tree_ = nullptr;   // whatever a tester left is overwritten

  StatScope stat(kStatTraversal, !is_replica_);
  unsigned int traversal_size = 0;
  int ok;
  if (type == TraversalType::FULL)
    ok = pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, cb_full_traversal,
                            travbuffer_.data(), &traversal_size);
  else if (type == TraversalType::PARTIAL)
    ok = pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, cb_partial_traversal,
                            travbuffer_.data(), &traversal_size);
  else
    throw std::invalid_argument("Invalid traversal type");
  if (!ok) throw std::invalid_argument("TraversalUpdate() requires an inner node");

  unsigned int matrix_count = 0, ops_count = 0;
  pll_utree_create_operations(travbuffer_.data(), traversal_size, branch_lengths_.data(),
                              matrix_indices_.data(), operations_.data(), &matrix_count,
                              &ops_count);
  OnShards([&](ptgpu_engine* e, unsigned int) {
    const int rc = ptgpu_update_prob_matrices(e, matrix_indices_.data(), branch_lengths_.data(),
                                              matrix_count);
    if (rc != PTGPU_OK) return rc;
    return ptgpu_update_partials(e, reinterpret_cast<const ptgpu_operation_t*>(operations_.data()),
                                 ops_count);
  });
  // a FULL traversal leaves the exact CLVs and matrices of this tree, oriented at `root`; a
  // PARTIAL one that recomputed something leaves a mixture nobody can name
  if (type == TraversalType::FULL) {
    content_fp_ = FingerprintFrom(root);
    content_id_ = ExactContentId(content_fp_, root->node_index);
  } else if (ops_count)
    content_id_ = 0;
  return ops_count;
}

// PTPLL/pll_partition.cpp:395-406
void Partition::UpdateBranchLength(pll_unode_t* node, double length)
{
  node->length = length;
  node->back->length = length;
  content_id_ = 0;
  OnShards([&](ptgpu_engine* e, unsigned int) {
    return ptgpu_update_prob_matrices(e, &node->pmatrix_index, &node->length, 1);
  });
}

// PTPLL/pll_partition.cpp:408-488.  The sumtable, the <= MAX_ITER Newton steps
// with the reference's stopping rules and the final P-matrix update run as one
// device-resident sequence; only the optimised length comes back.
double Partition::OptimizeBranch(pll_unode_t* node)
{
  StatScope stat(kStatOptimizeBranch, !is_replica_);
  content_id_ = 0;
  pll_unode_t* child = node->back;
  double len = node->length;
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    double mine = node->length;
    const int rc = ptgpu_optimize_branch(e, node->clv_index, node->scaler_index, child->clv_index,
                                         child->scaler_index, node->pmatrix_index, node->length,
                                         MAX_ITER, EPSILON, map_mode_ ? kMapLambda : 0.0, &mine, nullptr);
    if (s == 0) len = mine;
    return rc;
  });
  node->length = len;
  node->back->length = len;
  return len;
}

// PTPLL/pll_partition.cpp:490-540
void Partition::OptimizeAllBranchesOnce(pll_unode_t* tree, bool use_pllmod)
{
  if (use_pllmod) {
    TraversalUpdate(tree, TraversalType::PARTIAL);
    LocalBranchOptimization(tree, 1e-8, 100.0, 1e-8, 1, -1);
    TraversalUpdate(tree, TraversalType::PARTIAL);
    return;
  }

  std::vector<pll_unode_t*> nodes(node_count(), nullptr);
  unsigned int found = 0;
  if (!pll_utree_traverse(tree, PLL_TREE_TRAVERSE_POSTORDER, visit_all, nodes.data(), &found))
    throw std::invalid_argument("OptimizeAllBranches() requires an inner node");
  if (found != nodes.size()) throw std::invalid_argument("Unexpected number of nodes");

  for (pll_unode_t* node : nodes) {
    if (!node->next) node = node->back;  // a tip's branch is optimised from its parent
    TraversalUpdate(node, TraversalType::PARTIAL);
    OptimizeBranch(node);
  }
}

// PTPLL/pll_partition.cpp:542-582
namespace {
std::function<void(Partition&, pll_utree_t*, pll_unode_t*, bool)> g_orientation_hook;
}

void Partition::SetOrientationHook(std::function<void(Partition&, pll_utree_t*, pll_unode_t*, bool)> hook)
{
  g_orientation_hook = std::move(hook);
}

void Partition::DeferTesterOrientation(pll_utree_t* tree, pll_unode_t* tested, bool invalidate_edge)
{
  if (is_replica_) return;
  pending_tree_ = tree;
  pending_node_ = tested;
  pending_invalidate_ = invalidate_edge;
  InvalidationLog().clear();   // what is invalidated from here on happens after the tester's last state
}

// see DeferTesterOrientation in the header
void Partition::ProduceTesterOrientation(pll_unode_t* visit_root)
{
  pll_utree_t* tree = pending_tree_;
  pll_unode_t* tested = pending_node_;
  const bool invalidate = pending_invalidate_;
  pending_tree_ = nullptr;
  const std::vector<pll_unode_t*> since = InvalidationLog();   // (copy: the calls below log too)
  TraversalUpdate(tested, TraversalType::FULL);
  if (invalidate) InvalidateEdgeClvs(tested);
  for (pll_unode_t* n : since) InvalidateEdgeClvs(n);
  // the tree being visited was cloned from `tree` (records keep their indices under an NNI) with the
  // flags it had then (src/wanderer.cpp:160-166): give it the ones it would have had
  const unsigned int records = tree->tip_count + 3 * tree->inner_count;
  std::vector<signed char> valid(records, -1);
  for (unsigned int i = 0; i < tree->tip_count + tree->inner_count; ++i) {
    pll_unode_t* first = tree->nodes[i];
    pll_unode_t* r = first;
    do {
      if (r->data && r->node_index < records) valid[r->node_index] = static_cast<node_info_t*>(r->data)->clv_valid ? 1 : 0;
      r = r->next;
    } while (r && r != first);
  }
  unsigned int found = 0;
  if (pll_utree_traverse(visit_root, PLL_TREE_TRAVERSE_POSTORDER, visit_all, travbuffer_.data(), &found) &&
      found == node_count()) {
    for (unsigned int i = 0; i < found; ++i) {
      pll_unode_t* first = travbuffer_[i];
      pll_unode_t* r = first;
      do {
        if (r->data && r->node_index < records && valid[r->node_index] >= 0)
          static_cast<node_info_t*>(r->data)->clv_valid = valid[r->node_index];
        r = r->next;
      } while (r && r != first);
    }
  }
  if (g_orientation_hook) g_orientation_hook(*this, tree, tested, invalidate);
}

void Partition::OptimizeAllBranches(pll_unode_t* tree, bool use_pllmod)
{
  StatScope stat(kStatAllBranches, !is_replica_);
  if (pending_tree_ && !use_pllmod && !is_replica_) ProduceTesterOrientation(tree);
  if (!use_pllmod && !is_replica_ && g_visits_used.load(std::memory_order_relaxed)) {
    const uint64_t key = VisitKey(tree);
    const bool hit = key && TakePublishedVisit(key, tree);
    static const bool debug = std::getenv("PTGPU_PREFETCH_DEBUG") != nullptr;
    if (debug)
      std::fprintf(stderr, "[prefetch] wanderer key %016llx content %016llx %s\n", (unsigned long long)key,
                   (unsigned long long)content_id_, hit ? "hit" : "MISS");
    if (hit) {
      g_visits_served += 1;
      return;
    }
  }
  if (!is_replica_) g_visits_computed += 1;
  // whichever way it is computed, the procedure ends with every CLV and matrix exact for the
  // optimised tree, oriented at `tree`
  struct ContentNote {
    Partition* p; pll_unode_t* root; bool armed;
    ~ContentNote()
    {
      const bool ok = armed && std::uncaught_exceptions() == 0;
      p->content_fp_ = ok ? p->FingerprintFrom(root) : 0;
      p->content_id_ = ok ? ExactContentId(p->content_fp_, root->node_index) : 0;
    }
  } note{this, tree, !use_pllmod};
  content_id_ = 0;
  if (!use_pllmod && !stepwise_all_branches_) {
    OptimizeAllBranchesFused(tree);
    return;
  }
  TraversalUpdate(tree, TraversalType::PARTIAL);

  if (use_pllmod) {
    LocalBranchOptimization(tree, 1e-8, 100.0, 1e-8, static_cast<int>(MAX_ITER), -1);
    TraversalUpdate(tree, TraversalType::PARTIAL);
    return;
  }

  double before = LogLikelihood(tree);
  OptimizeAllBranchesOnce(tree);
  TraversalUpdate(tree, TraversalType::PARTIAL);
  double after = LogLikelihood(tree);

  for (unsigned int i = 0; std::fabs(before - after) > EPSILON && i < MAX_ITER; ++i) {
    before = after;
    OptimizeAllBranchesOnce(tree);
    TraversalUpdate(tree, TraversalType::PARTIAL);
    after = LogLikelihood(tree);
  }
}

// PTPLL/pll_partition.cpp:584-638
void Partition::OptimizeBranchNeighborhood(pll_unode_t* node, int radius)
{
  if (radius <= 0) throw std::invalid_argument("optimization radius must be positive");
  content_id_ = 0;
  TraversalUpdate(node, TraversalType::PARTIAL);
  // PLLMOD_OPT_MIN_BRANCH_LEN, PLLMOD_OPT_MAX_BRANCH_LEN (pll_optimize.h:60-61)
  LocalBranchOptimization(node, 1.0e-4, 100.0, EPSILON, static_cast<int>(MAX_ITER), radius);
}

// One edge of pllmod's recomp_iterative (PLLMOD/optimize/pll_optimize.c:771-934,
// keep_update = 1, built with NOCHECK_PERBRANCH_IMPR => 30 Newton iterations):
// optimise the edge at `p` on the device, then, while the radius lasts, turn
// the CLV of p's node towards each of the two other edges, recurse, and turn
// it back.  false = the reference's PLL_FAILURE (derivative not finite or the
// iteration limit hit); the edge keeps its length in that case.
bool Partition::OptimizeEdgeBounded(pll_unode_t* p, double bl_min, double bl_max,
                                    double tolerance, int radius)
{
  pll_unode_t* q = p->next;
  pll_unode_t* z = q ? q->next : nullptr;

  double guess = p->length;
  if (guess < bl_min || guess > bl_max) guess = 0.1;  // PLLMOD_OPT_DEFAULT_BRANCH_LEN

  double len = guess;
  int status = 0;
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    double mine = guess;
    int mine_status = 0;
    const int rc = ptgpu_optimize_branch_bounded(e, p->clv_index, p->scaler_index, p->back->clv_index,
                                                 p->back->scaler_index, guess, bl_min, bl_max,
                                                 tolerance, 30, &mine, &mine_status, nullptr);
    if (s == 0) { len = mine; status = mine_status; }
    return rc;
  });
  if (status != 0) return false;

  if (std::fabs(p->length - len) > 1e-10)
    OnShards([&](ptgpu_engine* e, unsigned int) {
      return ptgpu_update_prob_matrices(e, &p->pmatrix_index, &len, 1);
    });
  p->length = len;
  p->back->length = len;

  if (radius && q && z) {
    auto reorient = [this](pll_unode_t* parent, pll_unode_t* c1, pll_unode_t* c2) {
      ptgpu_operation_t op;
      op.parent_clv_index = parent->clv_index;
      op.parent_scaler_index = parent->scaler_index;
      op.child1_clv_index = c1->back->clv_index;
      op.child1_matrix_index = c1->back->pmatrix_index;
      op.child1_scaler_index = c1->back->scaler_index;
      op.child2_clv_index = c2->back->clv_index;
      op.child2_matrix_index = c2->back->pmatrix_index;
      op.child2_scaler_index = c2->back->scaler_index;
      OnShards([&](ptgpu_engine* e, unsigned int) { return ptgpu_update_partials(e, &op, 1); });
    };
    reorient(q, p, z);
    if (!OptimizeEdgeBounded(q->back, bl_min, bl_max, tolerance, radius - 1)) return false;
    reorient(z, q, p);
    if (!OptimizeEdgeBounded(z->back, bl_min, bl_max, tolerance, radius - 1)) return false;
    reorient(p, z, q);
  }
  return true;
}

// pllmod_opt_optimize_branch_lengths_local (PLLMOD/optimize/pll_optimize.c:969-1091);
// radius < 0 = the whole tree (pllmod_opt_optimize_branch_lengths_iterative, :1107-1130).
// Returns what the reference returns: -lnL, or -0.0 after a failure.
double Partition::LocalBranchOptimization(pll_unode_t* tree, double bl_min, double bl_max,
                                          double tolerance, int smoothings, int radius)
{
  content_id_ = 0;   // edges are re-optimised (and their matrices rewritten) below
  auto edge_lnl = [this, tree]() {
    double v = 0.0;
    OnShards([&](ptgpu_engine* e, unsigned int s) {
      double mine = 0.0;
      const int rc = ptgpu_compute_edge_loglikelihood(e, tree->back->clv_index, tree->back->scaler_index,
                                                      tree->clv_index, tree->scaler_index,
                                                      tree->pmatrix_index, nullptr, &mine);
      if (s == 0) v = mine;
      return rc;
    });
    return v;
  };
  double lnl = edge_lnl();

  const double xmin = bl_min > 0 ? bl_min : 1.0e-4;
  const double xmax = bl_max > 0 ? bl_max : 100.0;
  const double xtol = bl_min > 0 ? bl_min / 10.0 : 1.0e-4;

  unsigned int iters = static_cast<unsigned int>(smoothings);
  while (iters) {
    if (!OptimizeEdgeBounded(tree, xmin, xmax, xtol, radius) ||
        !OptimizeEdgeBounded(tree->back, xmin, xmax, xtol, radius - 1)) {
      lnl = 0.0;  // PLL_FAILURE
      break;
    }
    const double now = edge_lnl();
    if (!(now - lnl > now * 1e-14)) {
      lnl = now;
      break;
    }
    --iters;
    if (std::fabs(now - lnl) < tolerance) iters = 0;
    lnl = now;
  }
  return -1 * lnl;
}

// The same procedure as one persistent kernel (ptgpu_optimize_all_branches).  Which
// matrices and CLVs each PARTIAL traversal recomputes depends only on the orientation
// flags, which live on the host: the traversals are run here for their flag updates
// and their operation lists are RECORDED instead of launched; the device then walks
// the recorded steps, does the Newton solves and decides after every pass whether
// another one is needed.  The engine ends in the state the call sequence would have
// left (same buffers written by the same operations), the tree gets the optimised
// lengths, the flags are those of the last traversal.
void Partition::OptimizeAllBranchesFused(pll_unode_t* tree)
{
  std::vector<pll_unode_t*> nodes(node_count(), nullptr);
  unsigned int found = 0;
  if (!pll_utree_traverse(tree, PLL_TREE_TRAVERSE_POSTORDER, visit_all, nodes.data(), &found))
    throw std::invalid_argument("OptimizeAllBranches() requires an inner node");
  if (found != nodes.size()) throw std::invalid_argument("Unexpected number of nodes");

  // orientation flags of every inner-node record, for the fix-point test below
  auto flags = [&nodes]() {
    std::vector<int> f;
    for (pll_unode_t* n : nodes) {
      if (!n->next) continue;
      pll_unode_t* r = n;
      do {
        f.push_back(r->data ? static_cast<node_info_t*>(r->data)->clv_valid : -1);
        r = r->next;
      } while (r != n);
    }
    return f;
  };
  auto restore = [&nodes](const std::vector<int>& f) {
    size_t k = 0;
    for (pll_unode_t* n : nodes) {
      if (!n->next) continue;
      pll_unode_t* r = n;
      do {
        if (r->data) static_cast<node_info_t*>(r->data)->clv_valid = f[k];
        ++k;
        r = r->next;
      } while (r != n);
    }
  };

  // fresh[i]: the engine's matrix i is known to be P(current length of edge i), so a
  // traversal's recomputation of it would write the same bits and is left out
  std::vector<char> fresh(branch_count(), 0);
  std::vector<ptgpu_blo_step_t> current;
  auto record_traversal = [&](pll_unode_t* root) {
    unsigned int size = 0, matrix_count = 0, ops_count = 0;
    if (!pll_utree_traverse(root, PLL_TREE_TRAVERSE_POSTORDER, cb_partial_traversal,
                            travbuffer_.data(), &size))
      throw std::invalid_argument("TraversalUpdate() requires an inner node");
    pll_utree_create_operations(travbuffer_.data(), size, branch_lengths_.data(),
                                matrix_indices_.data(), operations_.data(), &matrix_count,
                                &ops_count);
    for (unsigned int k = 0; k < matrix_count; ++k) {
      const unsigned int m = matrix_indices_[k];
      if (fresh[m]) continue;
      fresh[m] = 1;
      ptgpu_blo_step_t st;
      std::memset(&st, 0, sizeof st);
      st.type = PTGPU_BLO_MAT;
      st.b_matrix = st.length_slot = m;
      current.push_back(st);
    }
    for (unsigned int k = 0; k < ops_count; ++k) {
      const pll_operation_t& op = operations_[k];
      ptgpu_blo_step_t st;
      std::memset(&st, 0, sizeof st);
      st.type = PTGPU_BLO_OP;
      st.a_clv = op.parent_clv_index; st.a_scaler = op.parent_scaler_index;
      st.b_clv = op.child1_clv_index; st.b_matrix = op.child1_matrix_index;
      st.b_scaler = op.child1_scaler_index;
      st.c_clv = op.child2_clv_index; st.c_matrix = op.child2_matrix_index;
      st.c_scaler = op.child2_scaler_index;
      current.push_back(st);
    }
  };
  auto record_pass = [&]() {  // OptimizeAllBranchesOnce + the traversal back to `tree`
    for (pll_unode_t* node : nodes) {
      if (!node->next) node = node->back;
      record_traversal(node);
      ptgpu_blo_step_t st;
      std::memset(&st, 0, sizeof st);
      st.type = PTGPU_BLO_OPT;
      st.a_clv = node->clv_index; st.a_scaler = node->scaler_index;
      st.b_clv = node->back->clv_index; st.b_scaler = node->back->scaler_index;
      st.b_matrix = st.length_slot = node->pmatrix_index;
      current.push_back(st);
      fresh[node->pmatrix_index] = 1;
    }
    record_traversal(tree);
  };

  record_traversal(tree);
  std::vector<ptgpu_blo_step_t> prologue;
  prologue.swap(current);

  // pass k's traversals start from the flags pass k-1 left; record passes until that
  // state repeats (in practice the second pass already does)
  const unsigned int max_passes = MAX_ITER + 1;
  std::vector<std::vector<ptgpu_blo_step_t>> bodies;
  std::vector<std::vector<int>> flags_after;
  std::vector<std::vector<char>> fresh_after;
  while (bodies.size() < max_passes) {
    record_pass();
    bodies.emplace_back();
    bodies.back().swap(current);
    flags_after.push_back(flags());
    fresh_after.push_back(fresh);
    const size_t k = bodies.size();
    if (k >= 2 && flags_after[k - 1] == flags_after[k - 2] && fresh_after[k - 1] == fresh_after[k - 2])
      break;
  }
  size_t n_body = 0;
  for (const auto& b : bodies) n_body = std::max(n_body, b.size());
  std::vector<ptgpu_blo_step_t> steps(prologue);
  for (const auto& b : bodies) {
    steps.insert(steps.end(), b.begin(), b.end());
    ptgpu_blo_step_t nop;
    std::memset(&nop, 0, sizeof nop);
    steps.insert(steps.end(), n_body - b.size(), nop);
  }

  // lengths by matrix index; LogLikelihood's edge and, in MAP mode, its prior terms
  std::vector<double> lengths(branch_count(), 0.0), optimised(branch_count(), 0.0);
  std::vector<unsigned int> prior;
  for (pll_unode_t* n : nodes) {
    lengths[n->pmatrix_index] = n->length;
    if (n->next) {
      lengths[n->next->pmatrix_index] = n->next->length;
      lengths[n->next->next->pmatrix_index] = n->next->next->length;
    }
    if (map_mode_ && n != tree->back) prior.push_back(n->pmatrix_index);
  }
  ptgpu_blo_edge_t edge{tree->clv_index, tree->scaler_index, tree->back->clv_index,
                        tree->back->scaler_index, tree->pmatrix_index, 0};
  double lnl = 0.0;
  int status = 0, passes = 0;
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    // every shard runs the same program and ends with the same lengths; shard 0's are taken
    std::vector<double> mine(s == 0 ? 0 : branch_count());
    double mine_lnl = 0.0;
    int mine_status = 0, mine_passes = 0;
    const int rc = ptgpu_optimize_all_branches(
        e, static_cast<unsigned int>(prologue.size()), static_cast<unsigned int>(n_body),
        static_cast<unsigned int>(bodies.size()), steps.data(), &edge, lengths.data(), prior.data(),
        static_cast<unsigned int>(prior.size()), MAX_ITER, EPSILON, map_mode_ ? kMapLambda : 0.0,
        max_passes, s == 0 ? optimised.data() : mine.data(), &mine_lnl, &mine_status, &mine_passes);
    if (s == 0) { lnl = mine_lnl; status = mine_status; passes = mine_passes; }
    return rc;
  });

  for (pll_unode_t* n : nodes) {
    pll_unode_t* r = n;
    do {
      r->length = optimised[r->pmatrix_index];
      r = r->next;
    } while (r && r != n);
  }
  if (passes >= 1 && static_cast<size_t>(passes) < bodies.size()) restore(flags_after[passes - 1]);
}

// Model-parameter optimisation is pll-modules' Brent / L-BFGS-B machinery
// driving libpll's CPU likelihood (out of scope;
// --optimize-models is off in every BASELINE config).
double Partition::OptimizeModelOnce(pll_unode_t*)
{
  throw std::logic_error("OptimizeModelOnce is not provided by the B200 engine build "
                         "(pll-modules model optimisation, out of scope)");
}

double Partition::OptimizeModel(pll_unode_t*)
{
  throw std::logic_error("OptimizeModel is not provided by the B200 engine build "
                         "(pll-modules model optimisation, out of scope)");
}

double Partition::OptimizeAllBranchesAndModel(pll_unode_t*, bool)
{
  throw std::logic_error("OptimizeAllBranchesAndModel is not provided by the B200 engine build "
                         "(pll-modules model optimisation, out of scope)");
}

// ---------------------------------------------------------------------------
// Batched NNI neighbourhood (the north-star path): all candidates of one
// Wanderer::QueueMoves in one engine call.
// ---------------------------------------------------------------------------
// directional buffer of an inner-node record: the CLV of its node seen from
// record->back, kept for all three records of every inner node
unsigned int Partition::DirectionalClv(const pll_unode_t* r) const
{
  return r->next ? tip_node_count_ + directional_base_ + (r->node_index - tip_node_count_)
                 : r->clv_index;
}

int Partition::DirectionalScaler(const pll_unode_t* r) const
{
  return r->next ? static_cast<int>(directional_base_ + (r->node_index - tip_node_count_)) : -1;
}

// All P matrices at the tree's current lengths and the directional CLVs listed in
// `needed` (engine CLV indices; anything that is not a directional buffer is ignored) plus
// the ones they are computed from -- one engine call each.  An NNI batch never reads a CLV
// that faces a tip (a third of the 3(n-2)) nor the one facing a cherry from its only inner
// neighbour, so those are not computed.
void Partition::UpdateDirectionalClvs(pll_utree_t* tree, const std::vector<unsigned int>& needed)
{
  if (!directional_base_)
    throw std::logic_error("directional CLV buffers were disabled (PTGPU_NO_DIRECTIONAL)");
  const unsigned int tips = tip_node_count_;
  const unsigned int inner = inner_node_count();
  auto dir_clv = [&](const pll_unode_t* r) { return DirectionalClv(r); };
  auto dir_scaler = [&](const pll_unode_t* r) { return DirectionalScaler(r); };

  // all P matrices at the tree's current lengths
  std::vector<unsigned int> midx;
  std::vector<double> mlen;
  std::vector<char> seen(branch_count(), 0);
  for (unsigned int i = 0; i < tips + inner; ++i) {
    pll_unode_t* n = tree->nodes[i];
    pll_unode_t* ring[3] = {n, n->next, n->next ? n->next->next : nullptr};
    for (pll_unode_t* r : ring) {
      if (!r || r->pmatrix_index >= seen.size() || seen[r->pmatrix_index]) continue;
      seen[r->pmatrix_index] = 1;
      midx.push_back(r->pmatrix_index);
      mlen.push_back(r->length);
    }
  }
  // every matrix becomes P(this tree's length of that edge): the same bits as before if the
  // content is this very tree's, something else otherwise
  if (content_id_ && TreeFingerprint(tree) != content_fp_) content_id_ = 0;
  OnShards([&](ptgpu_engine* e, unsigned int) {
    return ptgpu_update_prob_matrices(e, midx.data(), mlen.data(), static_cast<unsigned int>(midx.size()));
  });

  // one op per inner-node record, children before parents (iterative post-order)
  std::vector<ptgpu_operation_t> plan;
  plan.reserve(3 * inner);
  std::vector<char> done(3 * inner, 0);
  std::vector<std::pair<pll_unode_t*, bool>> stack;
  auto schedule = [&](pll_unode_t* start) {
    stack.emplace_back(start, false);
    while (!stack.empty()) {
      auto [r, expanded] = stack.back();
      stack.pop_back();
      if (!r->next || done[r->node_index - tips]) continue;
      pll_unode_t* a = r->next->back;
      pll_unode_t* b = r->next->next->back;
      if (!expanded) {
        stack.emplace_back(r, true);
        stack.emplace_back(b, false);
        stack.emplace_back(a, false);
        continue;
      }
      done[r->node_index - tips] = 1;
      ptgpu_operation_t op;
      op.parent_clv_index = dir_clv(r);
      op.parent_scaler_index = dir_scaler(r);
      op.child1_clv_index = dir_clv(a);
      op.child1_matrix_index = a->pmatrix_index;
      op.child1_scaler_index = dir_scaler(a);
      op.child2_clv_index = dir_clv(b);
      op.child2_matrix_index = b->pmatrix_index;
      op.child2_scaler_index = dir_scaler(b);
      plan.push_back(op);
    }
  };
  std::vector<char> wanted(3 * inner, 0);
  const unsigned int first = tips + directional_base_;
  for (unsigned int clv : needed)
    if (clv >= first && clv - first < wanted.size()) wanted[clv - first] = 1;
  for (unsigned int i = tips; i < tips + inner; ++i) {
    pll_unode_t* n = tree->nodes[i];
    pll_unode_t* ring[3] = {n, n->next, n->next->next};
    for (pll_unode_t* r : ring)
      if (wanted[r->node_index - tips]) schedule(r);
  }
  if (!plan.empty())
    OnShards([&](ptgpu_engine* e, unsigned int) {
      return ptgpu_update_partials(e, plan.data(), static_cast<unsigned int>(plan.size()));
    });
}

std::vector<Partition::NniScore> Partition::ScoreNniNeighborhood(pll_utree_t* tree, int radius)
{
  if (radius < 0) throw std::invalid_argument("optimization radius must not be negative");
  StatScope stat(kStatNeighborhood, !is_replica_);
  if (radius > 0) return ScoreNniNeighborhoodLocal(tree, radius);

  const unsigned int tips = tip_node_count_;
  const unsigned int inner = inner_node_count();
  auto dir_clv = [&](const pll_unode_t* r) { return DirectionalClv(r); };
  auto dir_scaler = [&](const pll_unode_t* r) { return DirectionalScaler(r); };

  // candidates in QueueMoves order (src/wanderer.cpp:242-285); after
  // pll_utree_nni(p, type) (LIBPLL/utree_moves.c:72-109) p's node joins
  // (y, w) and q's node joins (x, z)
  std::vector<NniScore> out;
  std::vector<ptgpu_nni_candidate_t> cands;
  auto child_of = [&](ptgpu_operation_t& op, int which, const pll_unode_t* c) {
    if (which == 0) {
      op.child1_clv_index = dir_clv(c);
      op.child1_matrix_index = c->pmatrix_index;
      op.child1_scaler_index = dir_scaler(c);
    } else {
      op.child2_clv_index = dir_clv(c);
      op.child2_matrix_index = c->pmatrix_index;
      op.child2_scaler_index = dir_scaler(c);
    }
  };
  for (unsigned int i = tips; i < tips + inner; ++i) {
    pll_unode_t* p = tree->nodes[i];
    pll_unode_t* q = p->back;
    if (!q->next) continue;
    for (int type : {PLL_UTREE_MOVE_NNI_LEFT, PLL_UTREE_MOVE_NNI_RIGHT}) {
      const pll_unode_t* x = p->next->back;
      const pll_unode_t* w = p->next->next->back;
      const pll_unode_t* y = type == PLL_UTREE_MOVE_NNI_LEFT ? q->next->back : q->next->next->back;
      const pll_unode_t* z = type == PLL_UTREE_MOVE_NNI_LEFT ? q->next->next->back : q->next->back;
      ptgpu_nni_candidate_t c;
      std::memset(&c, 0, sizeof c);
      c.p.parent_scaler_index = c.q.parent_scaler_index = -1;
      child_of(c.p, 0, y);
      child_of(c.p, 1, w);
      child_of(c.q, 0, x);
      child_of(c.q, 1, z);
      c.length0 = p->length;
      cands.push_back(c);
      out.push_back(NniScore{p, type, p->length, 0.0});
    }
  }
  if (cands.empty()) return out;

  // bring the directional CLVs the candidates read up to date
  std::vector<unsigned int> needed;
  needed.reserve(4 * cands.size());
  for (const ptgpu_nni_candidate_t& c : cands)
    for (const ptgpu_operation_t* op : {&c.p, &c.q}) {
      needed.push_back(op->child1_clv_index);
      needed.push_back(op->child2_clv_index);
    }
  UpdateDirectionalClvs(tree, needed);

  std::vector<double> len(cands.size()), lnl(cands.size());
  std::vector<int> status(cands.size());
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    std::vector<double> l2(s == 0 ? 0 : cands.size()), v2(s == 0 ? 0 : cands.size());
    std::vector<int> s2(s == 0 ? 0 : cands.size());
    return ptgpu_score_nni_batch(e, cands.data(), static_cast<unsigned int>(cands.size()), MAX_ITER,
                                 EPSILON, map_mode_ ? kMapLambda : 0.0, s == 0 ? len.data() : l2.data(),
                                 s == 0 ? lnl.data() : v2.data(), s == 0 ? status.data() : s2.data());
  });
  double prior = 0.0;
  if (map_mode_) prior = MapLogPrior(GetVirtualRoot(tree));
  for (size_t i = 0; i < out.size(); ++i) {
    if (status[i] != PTGPU_OK) throw std::range_error("d1 or d2 is not finite");
    out[i].length = len[i];
    out[i].lnl = lnl[i];
    if (map_mode_)  // swap the central branch's prior term for the one at its new length
      out[i].lnl += prior + kMapLambda * out[i].node->length - kMapLambda * len[i];
  }
  return out;
}

// ---------------------------------------------------------------------------
// radius >= 1: every candidate through pll-modules' local branch-length
// optimisation, all candidates in one persistent kernel
// (ptgpu_optimize_local_batch).  The host's part is integer bookkeeping: unroll
// recomp_iterative's recursion (PLLMOD/optimize/pll_optimize.c:771-934) around the
// moved edge into a flat step list of one common shape.
// ---------------------------------------------------------------------------
namespace {

struct ClvRef { bool is_private; unsigned int index; int scaler; };
struct MatRef { bool is_private; unsigned int index; };
struct SymStep {
  int type;
  ClvRef a, b, c;
  MatRef bm, cm;
  unsigned int slot;
};

// steps of recomp_iterative(., radius): one optimisation, and per level three
// re-orientations around two recursive calls
size_t RecompSize(int radius)
{
  size_t n = 1;
  for (int r = 0; r < radius; ++r) n = 4 + 2 * n;
  return n;
}

void RecompTypes(int radius, std::vector<int>& out)
{
  out.push_back(PTGPU_BLO_OPT);
  if (!radius) return;
  out.push_back(PTGPU_BLO_OP);
  RecompTypes(radius - 1, out);
  out.push_back(PTGPU_BLO_OP);
  RecompTypes(radius - 1, out);
  out.push_back(PTGPU_BLO_OP);
}

// symbolic program of one candidate; the reference keeps ONE CLV buffer per node,
// so "the CLV of a record's node" is the candidate's private copy once the
// recursion has written it, and the engine's directional CLV before that
class CandidateProgram {
 public:
  CandidateProgram(const Partition& part,
                   unsigned int (Partition::*clv)(const pll_unode_t*) const,
                   int (Partition::*scaler)(const pll_unode_t*) const)
      : part_(part), dir_clv_(clv), dir_scaler_(scaler) {}

  void ForgetSteps() { steps.clear(); private_nodes_.clear(); }

  ClvRef Clv(const pll_unode_t* r) const
  {
    if (!r->next) return ClvRef{false, r->clv_index, -1};
    for (size_t j = 0; j < private_nodes_.size(); ++j)
      if (private_nodes_[j] == r->clv_index) return ClvRef{true, static_cast<unsigned int>(j), 0};
    return ClvRef{false, (part_.*dir_clv_)(r), (part_.*dir_scaler_)(r)};
  }

  ClvRef Own(const pll_unode_t* r)
  {
    for (size_t j = 0; j < private_nodes_.size(); ++j)
      if (private_nodes_[j] == r->clv_index) return ClvRef{true, static_cast<unsigned int>(j), 0};
    private_nodes_.push_back(r->clv_index);
    return ClvRef{true, static_cast<unsigned int>(private_nodes_.size() - 1), 0};
  }

  MatRef Matrix(const pll_unode_t* r) const
  {
    for (size_t k = 0; k < slot_edges.size(); ++k)
      if (slot_edges[k] == r->pmatrix_index) return MatRef{true, static_cast<unsigned int>(k)};
    return MatRef{false, r->pmatrix_index};
  }

  unsigned int Slot(const pll_unode_t* r)
  {
    for (size_t k = 0; k < slot_edges.size(); ++k)
      if (slot_edges[k] == r->pmatrix_index) return static_cast<unsigned int>(k);
    slot_edges.push_back(r->pmatrix_index);
    lengths.push_back(r->length);
    return static_cast<unsigned int>(slot_edges.size() - 1);
  }

  // update_partials_and_scalers(parent, right, left), pll_optimize.c:741-768
  void Op(const pll_unode_t* parent, const pll_unode_t* c1, const pll_unode_t* c2)
  {
    SymStep st{};
    st.type = PTGPU_BLO_OP;
    st.a = Own(parent);
    st.b = Clv(c1->back);
    st.bm = Matrix(c1->back);
    st.c = Clv(c2->back);
    st.cm = Matrix(c2->back);
    steps.push_back(st);
  }

  void Opt(const pll_unode_t* r)
  {
    SymStep st{};
    st.type = PTGPU_BLO_OPT;
    st.slot = Slot(r);
    st.a = Clv(r);
    st.b = Clv(r->back);
    st.bm = MatRef{true, st.slot};
    steps.push_back(st);
  }

  void Nop(size_t n)
  {
    SymStep st{};
    st.type = PTGPU_BLO_NOP;
    steps.insert(steps.end(), n, st);
  }

  void Recomp(const pll_unode_t* r, int radius)
  {
    const size_t size = RecompSize(radius);
    if (!r) { Nop(size); return; }
    Opt(r);
    if (!radius) return;
    const pll_unode_t* q = r->next;
    if (!q) { Nop(size - 1); return; }
    const pll_unode_t* z = q->next;
    Op(q, r, z);
    Recomp(q->back, radius - 1);
    Op(z, q, r);
    Recomp(z->back, radius - 1);
    Op(r, z, q);
  }

  size_t private_count() const { return private_nodes_.size(); }

  std::vector<SymStep> steps;
  std::vector<unsigned int> slot_edges;  // slot -> pmatrix_index of the edge
  std::vector<double> lengths;           // slot -> length before the optimisation
  ClvRef smooth_parent, smooth_child, final_parent, final_child;
  MatRef edge_matrix;

 private:
  const Partition& part_;
  unsigned int (Partition::*dir_clv_)(const pll_unode_t*) const;
  int (Partition::*dir_scaler_)(const pll_unode_t*) const;
  std::vector<unsigned int> private_nodes_;  // private buffer -> clv_index of the node
};

} // namespace

std::vector<Partition::NniScore> Partition::ScoreNniNeighborhoodLocal(pll_utree_t* tree, int radius)
{
  const unsigned int tips = tip_node_count_;
  const unsigned int inner = inner_node_count();

  std::vector<int> body;
  RecompTypes(radius, body);
  RecompTypes(radius - 1, body);
  std::vector<int> types{PTGPU_BLO_OP, PTGPU_BLO_OP};
  types.insert(types.end(), body.begin(), body.end());
  types.insert(types.end(), body.begin(), body.end());

  std::vector<NniScore> out;
  std::vector<CandidateProgram> progs;
  for (unsigned int i = tips; i < tips + inner; ++i) {
    pll_unode_t* p = tree->nodes[i];
    pll_unode_t* q = p->back;
    if (!q->next) continue;
    for (int type : {PLL_UTREE_MOVE_NNI_LEFT, PLL_UTREE_MOVE_NNI_RIGHT}) {
      // the program is read off the tree WITH the move applied; the move is its own
      // inverse (LIBPLL/utree_moves.c:60-109) and leaves the CLV flags alone
      pll_utree_nni(p, type, nullptr);
      CandidateProgram c(*this, &Partition::DirectionalClv, &Partition::DirectionalScaler);
      // dry run: which edges get a private length / matrix slot
      c.Recomp(p, radius);
      c.Recomp(q, radius - 1);
      c.ForgetSteps();
      // TraversalUpdate(node, PARTIAL) after InvalidateEdgeClvs: q's node, then p's
      c.Op(q, q->next, q->next->next);
      c.Op(p, p->next, p->next->next);
      // first smoothing: nodes not re-oriented yet are read from the directional
      // CLVs; later smoothings: from the private copies the first one left behind
      for (int pass = 0; pass < 2; ++pass) {
        c.Recomp(p, radius);
        c.Recomp(q, radius - 1);
      }
      if (c.steps.size() != types.size()) throw std::logic_error("local program shape mismatch");
      c.smooth_parent = c.final_child = c.Clv(q);
      c.smooth_child = c.final_parent = c.Clv(p);
      c.edge_matrix = c.Matrix(p);
      pll_utree_nni(p, type, nullptr);
      out.push_back(NniScore{p, type, p->length, 0.0});
      progs.push_back(std::move(c));
    }
  }
  if (progs.empty()) return out;

  // bring the directional CLVs the programs read up to date
  {
    std::vector<unsigned int> needed;
    auto note = [&needed](const ClvRef& r) {
      if (!r.is_private) needed.push_back(r.index);
    };
    for (const CandidateProgram& c : progs) {
      for (const SymStep& st : c.steps) {
        if (st.type == PTGPU_BLO_NOP) continue;
        note(st.a);
        note(st.b);
        if (st.type == PTGPU_BLO_OP) note(st.c);
      }
      note(c.smooth_parent);
      note(c.smooth_child);
    }
    UpdateDirectionalClvs(tree, needed);
  }

  const unsigned int count = static_cast<unsigned int>(progs.size());
  unsigned int n_private = 1, n_lengths = 1;
  for (const CandidateProgram& c : progs) {
    n_private = std::max<unsigned int>(n_private, static_cast<unsigned int>(c.private_count()));
    n_lengths = std::max<unsigned int>(n_lengths, static_cast<unsigned int>(c.lengths.size()));
  }
  // private index spaces (include/ptgpu.h): above the engine's own ranges
  const unsigned int clv_total = tips + inner + (directional_base_ ? 3 * inner : 0);
  const unsigned int scaler_total = inner + (directional_base_ ? 3 * inner : 0);
  const unsigned int matrix_total = branch_count();

  std::vector<ptgpu_blo_step_t> steps(types.size() * count);
  std::vector<ptgpu_blo_edge_t> smooth(count), final_edge(count);
  std::vector<double> lengths0(static_cast<size_t>(count) * n_lengths, 0.1);
  for (unsigned int i = 0; i < count; ++i) {
    const CandidateProgram& c = progs[i];
    auto clv = [&](const ClvRef& r) {
      return r.is_private ? clv_total + i * n_private + r.index : r.index;
    };
    auto scaler = [&](const ClvRef& r) {
      return r.is_private ? static_cast<int>(scaler_total + i * n_private + r.index) : r.scaler;
    };
    auto matrix = [&](const MatRef& m) {
      return m.is_private ? matrix_total + i * n_lengths + m.index : m.index;
    };
    for (size_t k = 0; k < c.lengths.size(); ++k) lengths0[i * n_lengths + k] = c.lengths[k];
    for (size_t k = 0; k < c.steps.size(); ++k) {
      const SymStep& st = c.steps[k];
      ptgpu_blo_step_t& d = steps[k * count + i];
      std::memset(&d, 0, sizeof d);
      d.type = st.type;
      if (st.type == PTGPU_BLO_NOP) continue;
      d.a_clv = clv(st.a); d.a_scaler = scaler(st.a);
      d.b_clv = clv(st.b); d.b_scaler = scaler(st.b); d.b_matrix = matrix(st.bm);
      if (st.type == PTGPU_BLO_OP) {
        d.c_clv = clv(st.c); d.c_scaler = scaler(st.c); d.c_matrix = matrix(st.cm);
      }
      d.length_slot = st.slot;
    }
    smooth[i] = ptgpu_blo_edge_t{clv(c.smooth_parent), scaler(c.smooth_parent), clv(c.smooth_child),
                                 scaler(c.smooth_child), matrix(c.edge_matrix), 0};
    final_edge[i] = ptgpu_blo_edge_t{clv(c.final_parent), scaler(c.final_parent), clv(c.final_child),
                                     scaler(c.final_child), matrix(c.edge_matrix), 0};
  }

  std::vector<double> lengths(lengths0.size()), lnl(count);
  std::vector<int> status(count);
  // PLLMOD_OPT_MIN/MAX_BRANCH_LEN, EPSILON, MAX_ITER: PTPLL/pll_partition.cpp:629-637
  OnShards([&](ptgpu_engine* e, unsigned int s) {
    std::vector<double> l2(s == 0 ? 0 : lengths0.size()), v2(s == 0 ? 0 : count);
    std::vector<int> s2(s == 0 ? 0 : count);
    return ptgpu_optimize_local_batch(e, count, n_private, n_lengths, 2,
                                      static_cast<unsigned int>(body.size()), types.data(), steps.data(),
                                      smooth.data(), final_edge.data(), lengths0.data(), 1.0e-4, 100.0,
                                      EPSILON, MAX_ITER, s == 0 ? lengths.data() : l2.data(),
                                      s == 0 ? lnl.data() : v2.data(), s == 0 ? status.data() : s2.data(),
                                      nullptr);
  });
  // PTGPU_DUMP_LOCAL_TESTS=<file>: every candidate whose optimisation ended with a failure status (and
  // every 997th other one; PTGPU_DUMP_LOCAL_EVERY=n: every nth) as "<status> <score> <newick of the moved tree, written from the move's node,
  // lengths with 17 digits>" -- to replay single cases on the reference (tests/golden/ds5_replay_local_tests.py)
  if (const char* path = std::getenv("PTGPU_DUMP_LOCAL_TESTS")) {
    static std::mutex dump_mutex;
    static unsigned long long seen = 0;
    std::lock_guard<std::mutex> lock(dump_mutex);
    FILE* f = nullptr;
    for (unsigned int i = 0; i < count; ++i) {
      static const unsigned long long every = [] {
        const char* env = std::getenv("PTGPU_DUMP_LOCAL_EVERY");
        return env ? std::max(1ull, std::strtoull(env, nullptr, 10)) : 997ull;
      }();
      if (status[i] == 0 && (++seen % every) != 0) continue;
      if (!f) f = std::fopen(path, "a");
      if (!f) break;
      pll_utree_t* moved = pll_utree_clone(tree);
      if (!moved) continue;
      pll_unode_t* at = nullptr;
      for (unsigned int k = 0; k < moved->tip_count + moved->inner_count && !at; ++k) {
        pll_unode_t* first = moved->nodes[k];
        pll_unode_t* r = first;
        do {
          if (r->node_index == out[i].node->node_index) { at = r; break; }
          r = r->next;
        } while (r && r != first);
      }
      if (at) {
        pll_utree_nni(at, out[i].type, nullptr);
        std::string text;
        std::function<void(const pll_unode_t*)> sub = [&](const pll_unode_t* r) {
          char len[40];
          std::snprintf(len, sizeof len, ":%.17g", r->length);
          if (!r->next) { text += r->label; text += len; return; }
          text += "(";
          sub(r->next->back);
          text += ",";
          sub(r->next->next->back);
          text += ")";
          text += len;
        };
        text += "(";
        sub(at->back);
        text += ",";
        sub(at->next->back);
        text += ",";
        sub(at->next->next->back);
        text += ");";
        std::fprintf(f, "%d %.17g %s\n", status[i], lnl[i], text.c_str());
      }
      pll_utree_destroy(moved, nullptr);
    }
    if (f) std::fclose(f);
  }
  double prior = 0.0;
  if (map_mode_) prior = MapLogPrior(GetVirtualRoot(tree));
  for (unsigned int i = 0; i < count; ++i) {
    out[i].length = lengths[i * n_lengths];  // slot 0 is the central branch
    out[i].lnl = lnl[i];
    if (map_mode_) {  // LogLikelihood's prior (PTPLL/pll_partition.cpp:307-325) at the new lengths
      double p = prior;
      for (size_t k = 0; k < progs[i].lengths.size(); ++k)
        p += kMapLambda * progs[i].lengths[k] - kMapLambda * lengths[i * n_lengths + k];
      out[i].lnl += p;
    }
  }
  return out;
}

} } // namespace pt::pll
