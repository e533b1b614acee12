// pt_b200/csrc/engine.cu -- the C ABI of include/ptgpu.h over the sm_100a kernels.
//
// One ptgpu_engine = one alignment resident in the HBM of one device, with the
// same index spaces as a pll_partition_t.  There is no CPU
// fallback anywhere in this file: every entry point either runs CUDA kernels
// or returns an error.
#include "../../include/ptgpu.h"

#include "gtr_model.hpp"
#include "kernels.cuh"
#include "blo.cuh"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace ptgpu;

static_assert(sizeof(ptgpu_operation_t) == sizeof(DevOp), "op layout");
static_assert(sizeof(ptgpu_nni_candidate_t) == sizeof(DevCand), "candidate layout");
static_assert(sizeof(ptgpu_blo_step_t) == sizeof(BloStep), "blo step layout");
static_assert(sizeof(ptgpu_blo_edge_t) == sizeof(BloEdge), "blo edge layout");

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...)
{
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

// Scratch that grows while the engine is in use comes from the stream-ordered allocator:
// cudaMalloc / cudaFree synchronise the whole device (and, with peer access enabled, touch
// the peers' page tables), which must not happen while a site shard's kernel on that device
// is waiting for a peer whose host thread would be the one blocked in the allocation.
#define PT_MALLOC(e, ptr, bytes) cudaMallocAsync((void **)(ptr), (bytes), (e)->stream)
#define PT_FREE(e, ptr) ((ptr) ? cudaFreeAsync((void *)(ptr), (e)->stream) : cudaSuccess)

#define CU(call)                                                                        \
  do                                                                                    \
  {                                                                                     \
    cudaError_t err__ = (call);                                                         \
    if (err__ != cudaSuccess)                                                           \
      return fail(PTGPU_ERR_CUDA, "%s failed: %s (%s:%d)", #call,                       \
                  cudaGetErrorString(err__), __FILE__, __LINE__);                       \
  } while (0)

// host staging area in pinned, device-mapped memory: small inputs go up and
// scalar results come back without extra cudaMemcpy calls
struct Staging
{
  static constexpr size_t kBytes = 1 << 20;
  unsigned char *host = nullptr;
  unsigned char *dev = nullptr;
};
// every staged scalar object has its own slot in the first 4 KiB
constexpr size_t kStgResult = 0;      // one double (edge lnL)
constexpr size_t kStgBad = 192;       // first illegal tip character
constexpr size_t kStgNewton = 256;    // one NewtonState
constexpr size_t kStgActive = 448;    // two counters of the split-phase Newton loop
constexpr size_t kStgMatrixIdx = 480; // one matrix index going up
constexpr size_t kStgShardErr = 512;  // sticky error flag of the shard exchange (written by kernels)
static_assert(sizeof(ptgpu::NewtonState) <= kStgActive - kStgNewton, "staging slot too small");

} // namespace

struct ptgpu_engine
{
  int device = 0;
  cudaStream_t stream = nullptr;
  unsigned int tips = 0, clv_buffers = 0, sites = 0, cats = 0, prob_matrices = 0,
               scale_buffers = 0;
  unsigned int ntiles = 0;
  unsigned int nheavy = 0; // tiles of the register-heavy CLV kernels (kHeavyTile sites)
  unsigned int chunk_tiles = 1, nchunks = 0; // Newton reduction units (functions of `sites` only)
  bool model_set = false;
  DevState dev{};
  ModelConst model{};
  uint8_t *d_tips = nullptr;
  uint8_t *d_raw = nullptr;         // raw characters of the tip being encoded
  uint8_t *d_raw_all = nullptr;     // ... of all tips (batched upload)
  size_t raw_all_cap = 0;
  unsigned int *d_tip_of = nullptr;
  unsigned long long *d_bad = nullptr;
  double *d_sumtable = nullptr;
  double *d_persite = nullptr;      // per-site lnL scratch (allocated on first use)
  double *d_partials = nullptr;     // [ntiles * 2] tile partials of single-edge reductions
  unsigned int *d_counter = nullptr;
  NewtonState *d_state = nullptr;   // single-branch Newton state
  unsigned int *d_idx1 = nullptr;   // one matrix index (optimize_branch)
  // batch scratch (grown on demand)
  double *d_bsum = nullptr;
  unsigned int *d_bsclr = nullptr;
  NewtonState *d_bstate = nullptr;
  double *d_bpartials = nullptr;
  DevGroup *d_bcands = nullptr;     // candidate groups of the current chunk
  size_t bcap = 0;
  // local-BLO batch scratch (grown on demand)
  double *d_lclv = nullptr; unsigned int *d_lscal = nullptr; double *d_lmat = nullptr;
  double *d_lsum = nullptr, *d_llens = nullptr, *d_lpart = nullptr, *d_ltile = nullptr;
  BloStep *d_lsteps = nullptr; BloEdge *d_ledges = nullptr; BloCand *d_lcand = nullptr;
  NewtonState *d_lnst = nullptr; int *d_ltypes = nullptr, *d_lfin = nullptr; unsigned int *d_lprior = nullptr;
  unsigned int blo_max_chunk = 0;          // env PTGPU_BLO_MAX_CHUNK (0 = as many as memory holds)
  unsigned int blo_cluster_work = 2 * 8;   // largest (candidates x tiles) run as one cluster; env PTGPU_BLO_CLUSTER_WORK
  size_t lcap_cand = 0, lcap_clvs = 0, lcap_lens = 0, lcap_steps = 0, lcap_types = 0, lcap_prior = 0;
  int blo_blocks_per_sm = 0;
  // op / matrix upload buffers
  DevOp *d_ops = nullptr;
  size_t ops_cap = 0;
  unsigned int *d_midx = nullptr;
  double *d_mlen = nullptr;
  size_t midx_cap = 0, mlen_cap = 0;
  Staging stg;
  // pinned host arena for the descriptor uploads / result downloads of the batch entry points
  // (grown on demand): a copy to or from PAGEABLE memory goes through the driver's shared
  // staging buffers and waits there for the stream, which serialises the engines of a
  // many-wanderer process to two kernels in flight
  unsigned char *h_pin = nullptr;
  size_t h_pin_cap = 0;
  std::vector<unsigned char *> h_pin_old;
  unsigned int *d_weights = nullptr; // pattern weights (ptgpu_set_pattern_weights), model.weights points here
  bool counted = false;             // included in g_live_engines
  int coop_blocks_per_sm = 0;
  int sm_count = 0;
  // site-shard mode (ptgpu_set_allreduce / ptgpu_set_shard)
  ptgpu_allreduce_fn allreduce = nullptr;
  void *allreduce_ctx = nullptr;
  unsigned long long vec_end = ~0ull;
  double *d_red = nullptr;          // [2 * red_cap] per-branch sums handed to the hook
  size_t red_cap = 0;
  // native exchange between site shards (ptgpu_comm_init / ptgpu_shard_connect): the kernels sum
  // per-tree scalars over peer-mapped mailboxes themselves (kernels.cuh, ShardComm)
  ShardComm comm{};
  unsigned long long *d_mailbox = nullptr;   // this shard's mailbox (peers write into it)
  unsigned long long *d_seq = nullptr;
  void *ipc_open[kMaxShards] = {};           // peers' mailboxes opened from IPC handles
  int same_device_shards = 1;                // shards of this alignment on this device (tests: several)
  unsigned int *d_active = nullptr;
  unsigned long long counters[4] = {0, 0, 0, 0};
  unsigned long long lnl_sweeps = 0, final_sweeps = 0;
  int partials_version = 4; // PTGPU_PARTIALS=3 selects the register-resident-matrix walker
  double lnl_margin = 100.0; // PTGPU_LNL_MARGIN
  int prepare_version = 2;  // PTGPU_PREPARE=1 selects the register-resident-matrix prepare kernel
  // batched NNI scoring keeps compact sumtables (13 instead of 16 doubles per site: the column of
  // the model's zero eigenvalue folded into one double per site, kernels.cuh); PTGPU_COMPACT_SUMTABLE=0
  // turns it off; zero_col = -1 if the model has no (numerically) zero eigenvalue
  bool compact_sumtable = true;
  int zero_col = -1;
  // optional per-kernel timing (ptgpu_set_profiling): CUDA events on e->stream
  bool profiling = false;
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> pending;
  std::vector<cudaEvent_t> event_pool;
  double kernel_ms[PTGPU_KERNEL_KINDS] = {};
  unsigned long long kernel_launches[PTGPU_KERNEL_KINDS] = {};
};

namespace {

// Engines alive on each device of this process.  A cooperative launch needs its whole grid
// resident at once, so a full-device grid has to wait until every other engine's kernel has
// drained (and holds up the ones behind it): with several engines on a device (one per
// wanderer) each takes its share of the SMs instead (at most PTGPU_COOP_SHARE_MAX ways, default
// 2: measured on the DS1 walk with 8 and 16 wanderers, halves beat both the full grid and
// smaller shares).  Results do not depend on the grid size (fixed reduction trees).
std::atomic<int> g_live_engines[64];

unsigned int coop_grid_share(const ptgpu_engine *e, unsigned int full)
{
  const int live = g_live_engines[e->device & 63].load(std::memory_order_relaxed);
  static const int max_div = [] { const char *v = getenv("PTGPU_COOP_SHARE_MAX"); return v ? std::max(1, atoi(v)) : 2; }();
  unsigned int div = (unsigned int)std::min(std::max(live, 1), max_div);
  // shards of one alignment that share a device wait for one another inside their kernels:
  // their cooperative grids must all be resident at the same time
  div = std::max(div, (unsigned int)e->same_device_shards);
  return std::max(std::min(full, 16u), full / div);
}

} // namespace

namespace {
// RAII bracket around one kernel launch: counts it and, when profiling is on,
// records a start/stop event pair on the engine's stream.
struct LaunchScope
{
  ptgpu_engine *e;
  int kind;
  cudaEvent_t a = nullptr, b = nullptr;
  static cudaEvent_t take(ptgpu_engine *e)
  {
    cudaEvent_t ev = nullptr;
    if (!e->event_pool.empty()) { ev = e->event_pool.back(); e->event_pool.pop_back(); }
    else cudaEventCreate(&ev);
    return ev;
  }
  LaunchScope(ptgpu_engine *eng, int k) : e(eng), kind(k)
  {
    e->counters[0]++;
    e->kernel_launches[kind]++;
    if (e->profiling) { a = take(e); b = take(e); cudaEventRecord(a, e->stream); }
  }
  ~LaunchScope()
  {
    if (a) { cudaEventRecord(b, e->stream); e->pending.push_back({kind, {a, b}}); }
  }
};
} // namespace

namespace {

int check_clv(const ptgpu_engine *e, unsigned int idx, const char *what)
{
  if (idx >= e->tips + e->clv_buffers)
    return fail(PTGPU_ERR_ARG, "%s: clv index %u out of range [0,%u)", what, idx,
                e->tips + e->clv_buffers);
  return PTGPU_OK;
}
int check_scaler(const ptgpu_engine *e, int idx, const char *what)
{
  if (idx != PTGPU_SCALE_BUFFER_NONE && (idx < 0 || (unsigned int)idx >= e->scale_buffers))
    return fail(PTGPU_ERR_ARG, "%s: scaler index %d out of range", what, idx);
  return PTGPU_OK;
}
int check_matrix(const ptgpu_engine *e, unsigned int idx, const char *what)
{
  if (idx >= e->prob_matrices)
    return fail(PTGPU_ERR_ARG, "%s: matrix index %u out of range [0,%u)", what, idx,
                e->prob_matrices);
  return PTGPU_OK;
}
int check_op(const ptgpu_engine *e, const ptgpu_operation_t &op, bool need_parent)
{
  int rc;
  if (need_parent)
  {
    if (op.parent_clv_index < e->tips)
      return fail(PTGPU_ERR_ARG, "operation writes tip clv %u", op.parent_clv_index);
    if ((rc = check_clv(e, op.parent_clv_index, "op.parent"))) return rc;
    if ((rc = check_scaler(e, op.parent_scaler_index, "op.parent"))) return rc;
  }
  if ((rc = check_clv(e, op.child1_clv_index, "op.child1"))) return rc;
  if ((rc = check_clv(e, op.child2_clv_index, "op.child2"))) return rc;
  if ((rc = check_scaler(e, op.child1_scaler_index, "op.child1"))) return rc;
  if ((rc = check_scaler(e, op.child2_scaler_index, "op.child2"))) return rc;
  if ((rc = check_matrix(e, op.child1_matrix_index, "op.child1"))) return rc;
  if ((rc = check_matrix(e, op.child2_matrix_index, "op.child2"))) return rc;
  return PTGPU_OK;
}

// dispatch on the compile-time category count
#define DISPATCH_CATS(e, CALL)                                                          \
  switch ((e)->cats)                                                                    \
  {                                                                                     \
    case 1: { constexpr int C = 1; CALL; break; }                                       \
    case 2: { constexpr int C = 2; CALL; break; }                                       \
    case 4: { constexpr int C = 4; CALL; break; }                                       \
    case 8: { constexpr int C = 8; CALL; break; }                                       \
    default: return fail(PTGPU_ERR_ARG, "unsupported rate category count %u", (e)->cats); \
  }

template <int C>
int coop_occupancy(int *out)
{
  int a = 0, b = 0;
  const int ring = (int)sweep_ring_bytes<C>();
  CU(cudaFuncSetAttribute(newton_kernel<C, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
  CU(cudaFuncSetAttribute(newton_kernel<C, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
  // the few-work-items variants (deep ring; one CTA per SM is all they need)
  constexpr int FS = few_work_stages<C>();
  const int deep = (int)ring_bytes<C>(FS);
  CU(cudaFuncSetAttribute(newton_kernel<C, false, FS>, cudaFuncAttributeMaxDynamicSharedMemorySize, deep));
  CU(cudaFuncSetAttribute(newton_kernel<C, true, FS>, cudaFuncAttributeMaxDynamicSharedMemorySize, deep));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, false, false, FS>, cudaFuncAttributeMaxDynamicSharedMemorySize, deep));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, true, false, FS>, cudaFuncAttributeMaxDynamicSharedMemorySize, deep));
  CU(cudaFuncSetAttribute(newton_sweep_kernel<C, true, true, FS>, cudaFuncAttributeMaxDynamicSharedMemorySize, deep));
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, newton_kernel<C, false>, kSitesPerPass * C, ring));
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, newton_kernel<C, true>, kSitesPerPass * C, ring));
  *out = std::min(a, b);
  return PTGPU_OK;
}

} // namespace

extern "C" {

const char *ptgpu_last_error(void) { return g_err.c_str(); }

int ptgpu_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int ptgpu_create(unsigned int tips, unsigned int clv_buffers, unsigned int sites,
                 unsigned int rate_cats, unsigned int prob_matrices,
                 unsigned int scale_buffers, int device, ptgpu_engine **out)
{
  if (!out) return fail(PTGPU_ERR_ARG, "ptgpu_create: out is null");
  *out = nullptr;
  if (tips < 3 || sites == 0 || clv_buffers == 0 || prob_matrices == 0)
    return fail(PTGPU_ERR_ARG, "ptgpu_create: empty partition (tips=%u sites=%u)", tips, sites);
  if (!(rate_cats == 1 || rate_cats == 2 || rate_cats == 4 || rate_cats == 8))
    return fail(PTGPU_ERR_ARG, "ptgpu_create: rate_cats must be 1, 2, 4 or 8 (got %u)", rate_cats);
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev)
    return fail(PTGPU_ERR_ARG, "ptgpu_create: device %d not present (%d devices)", device, ndev);
  CU(cudaSetDevice(device));

  ptgpu_engine *e = new ptgpu_engine;
  e->device = device;
  if (const char *pv = getenv("PTGPU_PARTIALS")) e->partials_version = atoi(pv);
  if (const char *pv = getenv("PTGPU_PREPARE")) e->prepare_version = atoi(pv);
  if (const char *pv = getenv("PTGPU_LNL_MARGIN")) e->lnl_margin = atof(pv);
  if (const char *pv = getenv("PTGPU_COMPACT_SUMTABLE")) e->compact_sumtable = atoi(pv) != 0;
  if (const char *pv = getenv("PTGPU_BLO_CLUSTER_WORK")) e->blo_cluster_work = (unsigned int)atoi(pv);
  if (const char *pv = getenv("PTGPU_BLO_MAX_CHUNK")) e->blo_max_chunk = (unsigned int)atoi(pv);
  e->tips = tips; e->clv_buffers = clv_buffers; e->sites = sites; e->cats = rate_cats;
  e->prob_matrices = prob_matrices; e->scale_buffers = scale_buffers;
  e->ntiles = (sites + kTileSites - 1) / kTileSites;
  e->nheavy = (sites + kHeavyTile - 1) / kHeavyTile;
  // Reduction units of every sum over sites: chunks of up to 8 tiles (a tile = 256 sites), about a
  // hundred of them on a mid-sized alignment and ~490 at 1M sites, so that ONE branch's sweep (B = 1:
  // OptimizeBranch, every Newton step of a whole-tree optimisation) still spreads over all SMs, while a
  // chunk stays long enough that the per-chunk reduction does not weigh on the batched sweeps
  // (measured: 1-tile chunks at 100k sites held the batched Newton kernel at 0.64-0.72 of the HBM
  // peak).  A function of `sites` only: results do not depend on launch geometry.
  e->chunk_tiles = std::min(8u, std::max(1u, e->ntiles / 96u));
  e->nchunks = (e->ntiles + e->chunk_tiles - 1) / e->chunk_tiles;
  const size_t clv_stride = (size_t)sites * rate_cats * 4;
#define CU_NEW(call)                                                                    \
  do { cudaError_t err__ = (call); if (err__ != cudaSuccess) { int rc__ = fail(PTGPU_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(err__)); ptgpu_destroy(e); return rc__; } } while (0)
  CU_NEW(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  const size_t tip_stride = ((size_t)sites + 15) & ~(size_t)15;      // rows start 16-byte aligned (TMA bulk)
  const size_t scaler_stride = ((size_t)sites + 3) & ~(size_t)3;
  CU_NEW(cudaMalloc(&e->d_tips, (size_t)tips * tip_stride));
  CU_NEW(cudaMemsetAsync(e->d_tips, 0, (size_t)tips * tip_stride, e->stream));
  CU_NEW(cudaMalloc(&e->dev.clv, clv_stride * clv_buffers * sizeof(double)));
  CU_NEW(cudaMalloc(&e->dev.scalers, (size_t)std::max(1u, scale_buffers) * scaler_stride * sizeof(unsigned int)));
  CU_NEW(cudaMemsetAsync(e->dev.scalers, 0, (size_t)std::max(1u, scale_buffers) * scaler_stride * sizeof(unsigned int), e->stream));
  CU_NEW(cudaMalloc(&e->dev.pmat, (size_t)prob_matrices * rate_cats * 16 * sizeof(double)));
  CU_NEW(cudaMemsetAsync(e->dev.pmat, 0, (size_t)prob_matrices * rate_cats * 16 * sizeof(double), e->stream));
  CU_NEW(cudaMalloc(&e->d_sumtable, clv_stride * sizeof(double)));
  CU_NEW(cudaMalloc(&e->d_partials, (size_t)e->ntiles * 3 * sizeof(double)));
  CU_NEW(cudaMalloc(&e->d_counter, sizeof(unsigned int)));
  CU_NEW(cudaMemsetAsync(e->d_counter, 0, sizeof(unsigned int), e->stream));
  CU_NEW(cudaMalloc(&e->d_state, sizeof(NewtonState)));
  CU_NEW(cudaMalloc(&e->d_idx1, sizeof(unsigned int)));
  CU_NEW(cudaHostAlloc(&e->stg.host, Staging::kBytes, cudaHostAllocMapped));
  CU_NEW(cudaHostGetDevicePointer((void **)&e->stg.dev, e->stg.host, 0));
  e->dev.tips = e->d_tips;
  e->dev.n_tips = tips;
  e->dev.sites = sites;
  e->dev.clv_stride = clv_stride;
  e->dev.tip_stride = tip_stride;
  e->dev.scaler_stride = scaler_stride;
  CU_NEW(cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, device));
  {
    int rc = PTGPU_OK, occ = 0;
    switch (rate_cats)
    {
      case 1: rc = coop_occupancy<1>(&occ); break;
      case 2: rc = coop_occupancy<2>(&occ); break;
      case 4: rc = coop_occupancy<4>(&occ); break;
      default: rc = coop_occupancy<8>(&occ); break;
    }
    if (rc) { ptgpu_destroy(e); return rc; }
    if (occ < 1) { ptgpu_destroy(e); return fail(PTGPU_ERR_CUDA, "newton kernel does not fit an SM"); }
    e->coop_blocks_per_sm = occ;
    cudaError_t ea = cudaSuccess;
#define PT_SET_SMEM(CC)                                                                                   \
  ea = cudaFuncSetAttribute(nni_prepare_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                            (int)sizeof(PrepSmem<CC>));                                                   \
  if (ea == cudaSuccess)                                                                                  \
    ea = cudaFuncSetAttribute(nni_prepare_kernel_v2<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                              (int)sizeof(PrepSmem2<CC>));
    switch (rate_cats)
    {
      case 1: { PT_SET_SMEM(1) break; }
      case 2: { PT_SET_SMEM(2) break; }
      case 4: { PT_SET_SMEM(4) break; }
      default: { PT_SET_SMEM(8) break; }
    }
#undef PT_SET_SMEM
    if (ea != cudaSuccess) { ptgpu_destroy(e); return fail(PTGPU_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(ea)); }
  }
  CU_NEW(cudaStreamSynchronize(e->stream));
#undef CU_NEW
  e->counted = true;
  g_live_engines[e->device & 63].fetch_add(1);
  *out = e;
  return PTGPU_OK;
}

void ptgpu_destroy(ptgpu_engine *e)
{
  if (!e) return;
  if (e->counted) g_live_engines[e->device & 63].fetch_sub(1);
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (void *p : e->ipc_open)
    if (p) cudaIpcCloseMemHandle(p);
  cudaFree(e->d_mailbox); cudaFree(e->d_seq);
  cudaFree(e->d_raw_all); cudaFree(e->d_tip_of); cudaFree(e->d_bad); cudaFree(e->d_weights);
  cudaFree(e->d_tips); cudaFree(e->d_raw); cudaFree(e->dev.clv); cudaFree(e->dev.scalers); cudaFree(e->dev.pmat);
  cudaFree(e->d_sumtable); cudaFree(e->d_persite); cudaFree(e->d_partials); cudaFree(e->d_counter);
  cudaFree(e->d_state); cudaFree(e->d_idx1); cudaFree(e->d_red); cudaFree(e->d_active);
  cudaFree(e->d_bsum); cudaFree(e->d_bsclr); cudaFree(e->d_bstate); cudaFree(e->d_bpartials);
  cudaFree(e->d_lclv); cudaFree(e->d_lscal); cudaFree(e->d_lmat); cudaFree(e->d_lsum); cudaFree(e->d_llens);
  cudaFree(e->d_lpart); cudaFree(e->d_ltile); cudaFree(e->d_lsteps); cudaFree(e->d_ledges); cudaFree(e->d_lcand);
  cudaFree(e->d_lnst); cudaFree(e->d_ltypes); cudaFree(e->d_lprior); cudaFree(e->d_lfin);
  cudaFree(e->d_bcands); cudaFree(e->d_ops); cudaFree(e->d_midx); cudaFree(e->d_mlen);
  for (auto &p : e->pending) { cudaEventDestroy(p.second.first); cudaEventDestroy(p.second.second); }
  for (auto ev : e->event_pool) cudaEventDestroy(ev);
  if (e->stg.host) cudaFreeHost(e->stg.host);
  if (e->h_pin) cudaFreeHost(e->h_pin);
  for (unsigned char *p : e->h_pin_old) cudaFreeHost(p);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int ptgpu_set_model(ptgpu_engine *e, const double freqs[4], const double subst[6],
                    const double *category_rates)
{
  if (!e || !freqs || !subst || !category_rates) return fail(PTGPU_ERR_ARG, "ptgpu_set_model: null argument");
  ModelConst &m = e->model;
  gtr_eigensystem(freqs, subst, m.evecs, m.inv_evecs, m.evals);
  for (int i = 0; i < 4; ++i) m.freqs[i] = freqs[i];
  for (int j = 0; j < 4; ++j)
    for (int k = 0; k < 4; ++k) m.tt_inv[j * 4 + k] = m.inv_evecs[k * 4 + j] * freqs[k];
  for (unsigned int c = 0; c < kMaxCats; ++c) m.rates[c] = c < e->cats ? category_rates[c] : 0.0;
  m.rate_weight = 1.0 / e->cats;
  m.log_scale_thr = std::log(1.0 / std::ldexp(1.0, 256));
  // the eigenvalue 0 of a reversible model comes out of the QL iteration as ~1e-17 (its exact
  // value is 0): exp(lambda r t) is 1.0 to within 1e-15 for every length an optimiser can reach
  {
    double big = 0.0;
    for (int j = 0; j < 4; ++j) big = std::max(big, std::fabs(m.evals[j]));
    e->zero_col = -1;
    int hits = 0;
    for (int j = 0; j < 4; ++j)
      if (std::fabs(m.evals[j]) <= 1e-12 * big) { e->zero_col = j; ++hits; }
    if (hits != 1) e->zero_col = -1;
  }
  e->model_set = true;
  return PTGPU_OK;
}

int ptgpu_gamma_rates_mean(double alpha, unsigned int cats, double *rates_out)
{
  if (!gamma_mean_rates(alpha, cats, rates_out))
    return fail(PTGPU_ERR_ARG, "ptgpu_gamma_rates_mean: alpha %g / categories %u out of range", alpha, cats);
  return PTGPU_OK;
}

int ptgpu_get_eigen(ptgpu_engine *e, double evecs[16], double inv_evecs[16], double evals[4])
{
  if (!e || !e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  memcpy(evecs, e->model.evecs, sizeof e->model.evecs);
  memcpy(inv_evecs, e->model.inv_evecs, sizeof e->model.inv_evecs);
  memcpy(evals, e->model.evals, sizeof e->model.evals);
  return PTGPU_OK;
}

int ptgpu_set_tip_states(ptgpu_engine *e, unsigned int tip_index, const char *sequence)
{
  if (!e || !sequence) return fail(PTGPU_ERR_ARG, "ptgpu_set_tip_states: null argument");
  if (tip_index >= e->tips) return fail(PTGPU_ERR_ARG, "tip index %u out of range", tip_index);
  static const TipLut lut = [] {
    TipLut t;
    for (int c = 0; c < 256; ++c) t.m[c] = (uint8_t)nucleotide_mask((unsigned char)c);
    t.m[0] = 0;
    return t;
  }();
  CU(cudaSetDevice(e->device));
  if (!e->d_raw) CU(PT_MALLOC(e, &e->d_raw, e->sites));
  // the characters are encoded on the device: one H2D copy of the raw bytes
  // (asynchronous when the caller's buffer is pinned) and one byte-map kernel
  unsigned int *bad_h = (unsigned int *)(e->stg.host + kStgBad);
  unsigned int *bad_d = (unsigned int *)(e->stg.dev + kStgBad);
  CU(cudaStreamSynchronize(e->stream));
  *bad_h = 0xffffffffu;
  CU(cudaMemcpyAsync(e->d_raw, sequence, e->sites, cudaMemcpyHostToDevice, e->stream));
  {
    LaunchScope ls(e, PTGPU_KERNEL_TIP_ENCODE);
    const unsigned int blocks = std::min<unsigned int>((e->sites + 255) / 256, 4u * (unsigned int)e->sm_count);
    tip_encode_kernel<<<blocks, 256, 0, e->stream>>>((const unsigned char *)e->d_raw,
                                                     e->d_tips + (size_t)tip_index * e->dev.tip_stride,
                                                     e->sites, lut, bad_d);
  }
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(e->stream));
  const unsigned int bad = *(volatile unsigned int *)bad_h;
  if (bad != 0xffffffffu)
  {
    const unsigned char ch = (unsigned char)sequence[bad];
    return fail(PTGPU_ERR_STATE, "Illegal state code in tip \"%c\" (site %u)", ch ? ch : '?', bad);
  }
  return PTGPU_OK;
}

namespace { int pin_reserve(ptgpu_engine *e, size_t bytes); }

int ptgpu_set_pattern_weights(ptgpu_engine *e, const unsigned int *weights)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_set_pattern_weights: null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  if (!weights)
  {
    e->model.weights = nullptr;
    return PTGPU_OK;
  }
  for (unsigned int i = 0; i < e->sites; ++i)
    if (!weights[i]) return fail(PTGPU_ERR_ARG, "ptgpu_set_pattern_weights: weight of pattern %u is zero", i);
  if (!e->d_weights) CU(PT_MALLOC(e, &e->d_weights, (size_t)e->sites * sizeof(unsigned int)));
  int rc;
  if ((rc = pin_reserve(e, (size_t)e->sites * sizeof(unsigned int)))) return rc;
  memcpy(e->h_pin, weights, (size_t)e->sites * sizeof(unsigned int));
  CU(cudaMemcpyAsync(e->d_weights, e->h_pin, (size_t)e->sites * sizeof(unsigned int), cudaMemcpyHostToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  e->model.weights = e->d_weights;
  return PTGPU_OK;
}

int ptgpu_copy_tip_states(ptgpu_engine *dst, ptgpu_engine *src)
{
  if (!dst || !src) return fail(PTGPU_ERR_ARG, "ptgpu_copy_tip_states: null engine");
  if (dst->tips != src->tips || dst->sites != src->sites || dst->device != src->device)
    return fail(PTGPU_ERR_ARG, "ptgpu_copy_tip_states: engines differ in tips, sites or device");
  CU(cudaSetDevice(dst->device));
  CU(cudaStreamSynchronize(src->stream));
  CU(cudaMemcpyAsync(dst->d_tips, src->d_tips, (size_t)src->tips * src->dev.tip_stride, cudaMemcpyDeviceToDevice,
                     dst->stream));
  if (src->model.weights)
  {
    if (!dst->d_weights) CU(PT_MALLOC(dst, &dst->d_weights, (size_t)dst->sites * sizeof(unsigned int)));
    CU(cudaMemcpyAsync(dst->d_weights, src->d_weights, (size_t)src->sites * sizeof(unsigned int),
                       cudaMemcpyDeviceToDevice, dst->stream));
    dst->model.weights = dst->d_weights;
  }
  else
    dst->model.weights = nullptr;
  CU(cudaStreamSynchronize(dst->stream));
  return PTGPU_OK;
}

int ptgpu_compress_site_patterns(char **sequences, unsigned int count, unsigned int *length,
                                 unsigned int *weights_out)
{
  if (!sequences || !length || !weights_out || !count)
    return fail(PTGPU_ERR_ARG, "ptgpu_compress_site_patterns: null argument");
  const size_t n = *length;
  if (!n) return PTGPU_OK;
  // state codes of the characters and, for the way back, the LAST character of every code
  // (LIBPLL/compress.c:166-170: inv_charmap is filled in ascending character order)
  unsigned char code[256], back[256] = {0};
  for (int c = 0; c < 256; ++c) code[c] = (unsigned char)nucleotide_mask((unsigned char)c);
  code[0] = 0;
  for (int c = 0; c < 256; ++c)
    if (code[c]) back[code[c]] = (unsigned char)c;
  // columns as rows (:214-219)
  std::vector<unsigned char> col(n * count);
  for (unsigned int j = 0; j < count; ++j)
  {
    if (!sequences[j]) return fail(PTGPU_ERR_ARG, "ptgpu_compress_site_patterns: sequence %u is null", j);
    for (size_t i = 0; i < n; ++i)
    {
      const unsigned char k = code[(unsigned char)sequences[j][i]];
      if (!k)
        return fail(PTGPU_ERR_STATE, "Illegal state code in sequence %u \"%c\" (site %zu)", j,
                    sequences[j][i] ? sequences[j][i] : '?', i);
      col[i * count + j] = k;
    }
  }
  // the reference sorts the columns (multikey quicksort, :222) and keeps the first of every run
  // of equal ones (:229-241): the outcome is the set of distinct columns in ascending order
  std::vector<size_t> order(n);
  for (size_t i = 0; i < n; ++i) order[i] = i;
  const unsigned char *base = col.data();
  std::sort(order.begin(), order.end(), [base, count](size_t a, size_t b) {
    return memcmp(base + a * count, base + b * count, count) < 0;
  });
  size_t m = 0;
  for (size_t i = 0; i < n; ++i)
  {
    const unsigned char *c = base + order[i] * count;
    if (i && memcmp(c, base + order[i - 1] * count, count) == 0)
    {
      ++weights_out[m - 1];
      continue;
    }
    weights_out[m] = 1;
    order[m] = order[i];   // m <= i: entries before i are no longer needed
    ++m;
  }
  for (unsigned int j = 0; j < count; ++j)
  {
    for (size_t i = 0; i < m; ++i) sequences[j][i] = (char)back[base[order[i] * count + j]];
    sequences[j][m] = 0;   // :248-249 (the input carries its terminator, so index m <= n exists)
  }
  *length = (unsigned int)m;
  return PTGPU_OK;
}

int ptgpu_set_tip_states_batch(ptgpu_engine *e, unsigned int count, const unsigned int *tip_indices,
                               const char *const *sequences)
{
  if (!e || !tip_indices || !sequences) return fail(PTGPU_ERR_ARG, "ptgpu_set_tip_states_batch: null argument");
  if (!count) return PTGPU_OK;
  if (count > e->tips) return fail(PTGPU_ERR_ARG, "more sequences (%u) than tips (%u)", count, e->tips);
  for (unsigned int i = 0; i < count; ++i)
  {
    if (tip_indices[i] >= e->tips) return fail(PTGPU_ERR_ARG, "tip index %u out of range", tip_indices[i]);
    if (!sequences[i]) return fail(PTGPU_ERR_ARG, "ptgpu_set_tip_states_batch: null sequence");
  }
  static const TipLut lut = [] {
    TipLut t;
    for (int c = 0; c < 256; ++c) t.m[c] = (uint8_t)nucleotide_mask((unsigned char)c);
    t.m[0] = 0;
    return t;
  }();
  CU(cudaSetDevice(e->device));
  const size_t S = e->sites;
  if (e->raw_all_cap < (size_t)count * S)
  {
    if (e->d_raw_all) CU(PT_FREE(e, e->d_raw_all));
    e->d_raw_all = nullptr;
    e->raw_all_cap = 0;
    CU(PT_MALLOC(e, &e->d_raw_all, (size_t)e->tips * S));
    e->raw_all_cap = (size_t)e->tips * S;
  }
  if (!e->d_tip_of) CU(PT_MALLOC(e, &e->d_tip_of, e->tips * sizeof(unsigned int)));
  if (!e->d_bad) CU(PT_MALLOC(e, &e->d_bad, sizeof(unsigned long long)));
  unsigned long long *bad_h = (unsigned long long *)(e->stg.host + kStgBad);
  unsigned long long *bad_d = e->d_bad;
  CU(cudaStreamSynchronize(e->stream));
  CU(cudaMemsetAsync(bad_d, 0xff, sizeof(unsigned long long), e->stream));
  CU(cudaMemcpyAsync(e->d_tip_of, tip_indices, count * sizeof(unsigned int), cudaMemcpyHostToDevice, e->stream));
  // runs of sequences that are adjacent in host memory go up as one copy
  for (unsigned int i = 0; i < count;)
  {
    unsigned int j = i + 1;
    while (j < count && sequences[j] == sequences[j - 1] + S) ++j;
    CU(cudaMemcpyAsync(e->d_raw_all + (size_t)i * S, sequences[i], (size_t)(j - i) * S, cudaMemcpyHostToDevice,
                       e->stream));
    i = j;
  }
  {
    LaunchScope ls(e, PTGPU_KERNEL_TIP_ENCODE);
    const unsigned long long n = (unsigned long long)count * S;
    const unsigned int blocks = (unsigned int)std::min<unsigned long long>((n + 255) / 256, 16ull * e->sm_count);
    tip_encode_batch_kernel<<<blocks, 256, 0, e->stream>>>((const unsigned char *)e->d_raw_all, e->d_tips,
                                                           e->d_tip_of, count, (unsigned int)S,
                                                           e->dev.tip_stride, lut, bad_d);
  }
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(bad_h, bad_d, sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  const unsigned long long bad = *(volatile unsigned long long *)bad_h;
  if (bad != ~0ull)
  {
    const unsigned int k = (unsigned int)(bad / S), site = (unsigned int)(bad % S);
    const unsigned char ch = (unsigned char)sequences[k][site];
    return fail(PTGPU_ERR_STATE, "Illegal state code in tip \"%c\" (tip %u, site %u)", ch ? ch : '?',
                tip_indices[k], site);
  }
  return PTGPU_OK;
}

} // extern "C"

namespace {

// staging layout: [0, 4096) scalar results written by kernels through the
// mapped pointer (slots: kStg* above); [4096, 1 MiB) small uploads (ops, matrix lists, candidates)
constexpr size_t kUploadOff = 4096;

template <typename T>
int ensure_on(ptgpu_engine *e, T **ptr, size_t *cap, size_t need)
{
  if (need <= *cap) return PTGPU_OK;
  if (*ptr) CU(PT_FREE(e, *ptr));
  *ptr = nullptr;
  *cap = 0;
  size_t n = std::max<size_t>(need, 64);
  CU(PT_MALLOC(e, ptr, n * sizeof(T)));
  *cap = n;
  return PTGPU_OK;
}
#define ensure(ptr, cap, need) ensure_on(e, ptr, cap, need)

// pinned host arena of at least `bytes` (contents are NOT kept when it grows; the stream is
// drained first, so nothing in flight still reads or writes the old one)
int pin_reserve(ptgpu_engine *e, size_t bytes)
{
  if (bytes <= e->h_pin_cap) return PTGPU_OK;
  CU(cudaStreamSynchronize(e->stream));
  if (e->h_pin) e->h_pin_old.push_back(e->h_pin);   // freed with the engine: cudaFreeHost synchronises the device
  e->h_pin = nullptr;
  e->h_pin_cap = 0;
  const size_t n = std::max<size_t>(bytes + bytes / 2, 1 << 16);
  CU(cudaHostAlloc(&e->h_pin, n, cudaHostAllocDefault));
  e->h_pin_cap = n;
  return PTGPU_OK;
}

// carves 256-byte aligned sections out of the pinned arena
struct PinCursor
{
  unsigned char *base;
  size_t off = 0;
  template <typename T>
  T *take(size_t n)
  {
    T *p = reinterpret_cast<T *>(base + off);
    off += (n * sizeof(T) + 255) & ~size_t(255);
    return p;
  }
  static size_t padded(size_t bytes) { return (bytes + 255) & ~size_t(255); }
};

// host -> device copy of a small descriptor array on the engine's stream.
// Arrays that fit go through the pinned staging area (truly asynchronous);
// the stream is drained first so the area is never overwritten in flight.
int upload(ptgpu_engine *e, void *dst, const void *src, size_t bytes)
{
  if (bytes <= Staging::kBytes - kUploadOff)
  {
    CU(cudaStreamSynchronize(e->stream));
    memcpy(e->stg.host + kUploadOff, src, bytes);
    CU(cudaMemcpyAsync(dst, e->stg.host + kUploadOff, bytes, cudaMemcpyHostToDevice, e->stream));
  }
  else
  {
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, e->stream));
    CU(cudaStreamSynchronize(e->stream));
  }
  return PTGPU_OK;
}

template <int C>
int launch_pmatrix(ptgpu_engine *e, const unsigned int *d_idx, const double *d_len, unsigned int count)
{
  const unsigned int threads = 128, blocks = (count * C + threads - 1) / threads;
  {
    LaunchScope ls(e, PTGPU_KERNEL_PMATRIX);
    pmatrix_kernel<C><<<blocks, threads, 0, e->stream>>>(e->dev, e->model, d_idx, d_len, count);
  }
  CU(cudaGetLastError());
  return PTGPU_OK;
}

// few work items, each a long chunk: the launches that take the deep-ring sweep (kernels.cuh,
// few_work_stages); PTGPU_FEW_WORK_RING=0 turns it off
bool few_long_items(const ptgpu_engine *e, unsigned int work, unsigned int chunk_tiles)
{
  static const bool on = [] { const char *v = getenv("PTGPU_FEW_WORK_RING"); return !v || atoi(v) != 0; }();
  return on && work <= (unsigned int)e->sm_count && chunk_tiles >= 8;
}

template <int C>
int launch_newton(ptgpu_engine *e, NewtonArgs &a)
{
  const unsigned int maxg = coop_grid_share(e, (unsigned int)(e->coop_blocks_per_sm * e->sm_count));
  // work items of ~32 tiles when there are plenty (a batch of candidates), single chunks when
  // one branch has to fill the device
  // (the kernel lowers the group size itself when few branches are still running)
  a.group = std::max(1u, std::min(32u / std::max(1u, a.chunk_tiles), a.B * a.nchunks / (8u * maxg)));
  const unsigned int work = std::max(1u, a.B * ((a.nchunks + a.group - 1) / a.group));
  const unsigned int grid = std::min(std::max(1u, a.B * a.nchunks), maxg);
  void *params[] = {&a};
  {
    LaunchScope ls(e, PTGPU_KERNEL_NEWTON);
    if (few_long_items(e, work, a.chunk_tiles * a.group))
    {
      constexpr int FS = few_work_stages<C>();
      void *fn = a.want_lnl ? (void *)newton_kernel<C, true, FS> : (void *)newton_kernel<C, false, FS>;
      CU(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kSitesPerPass * C), params, ring_bytes<C>(FS), e->stream));
    }
    else
    {
      void *fn = a.want_lnl ? (void *)newton_kernel<C, true> : (void *)newton_kernel<C, false>;
      CU(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kSitesPerPass * C), params, sweep_ring_bytes<C>(), e->stream));
    }
  }
  return PTGPU_OK;
}

// this engine holds a site shard: its sums over sites are partial sums
inline bool is_sharded(const ptgpu_engine *e) { return e->allreduce != nullptr || e->comm.nranks > 1; }

// sum `count` doubles at `buf` (device) over all shards: through the caller's hook if one is
// set, else through the engines' own peer-memory exchange (one single-CTA kernel on the stream)
int shard_sum(ptgpu_engine *e, double *buf, size_t count)
{
  if (e->allreduce)
  {
    const int rc = e->allreduce(e->allreduce_ctx, buf, count, (void *)e->stream);
    if (rc != 0) return fail(PTGPU_ERR_STATE, "allreduce hook failed with code %d", rc);
    return PTGPU_OK;
  }
  for (size_t off = 0; off < count; off += kShardSlots)
  {
    const unsigned int n = (unsigned int)std::min<size_t>(kShardSlots, count - off);
    e->counters[0]++;
    shard_allreduce_kernel<<<1, 256, 0, e->stream>>>(e->comm, buf + off, n);
  }
  CU(cudaGetLastError());
  return PTGPU_OK;
}

// after a stream synchronisation: did an exchange between the shards time out?
int shard_check(ptgpu_engine *e)
{
  if (e->comm.nranks > 1 && *(volatile int *)(e->stg.host + kStgShardErr))
    return fail(PTGPU_ERR_STATE, "site shards: a peer shard did not answer within %llu ms (rank %d of %d)",
                e->comm.timeout_ns / 1000000ull, e->comm.rank, e->comm.nranks);
  return PTGPU_OK;
}

// Newton over site shards: local sweep + fixed-order local reduction, the hook
// sums the [2B] buffer over shards, every shard applies the same scalar update
// (so all shards stay in lock step and leave the loop together).
template <int C>
int newton_split(ptgpu_engine *e, NewtonArgs &a)
{
  int rc;
  if ((rc = ensure(&e->d_red, &e->red_cap, (size_t)a.B * 3))) return rc;
  if (!e->d_active) CU(PT_MALLOC(e, &e->d_active, 2 * sizeof(unsigned int)));
  const unsigned int work = std::max(1u, a.B * a.nchunks);
  const unsigned int grid = std::min(work, (unsigned int)e->sm_count * 8u);
  constexpr int FS = few_work_stages<C>();
  const bool deep = few_long_items(e, work, a.chunk_tiles);
  volatile unsigned int *active_h = (volatile unsigned int *)(e->stg.host + kStgActive);
  unsigned int pending_lnl = 0;
  for (unsigned int it = 0; it < a.max_iter; ++it)
  {
    {
      LaunchScope ls(e, PTGPU_KERNEL_NEWTON);
      if (deep)
      {
        if (a.want_lnl) newton_sweep_kernel<C, true, false, FS><<<grid, kSitesPerPass * C, ring_bytes<C>(FS), e->stream>>>(a);
        else newton_sweep_kernel<C, false, false, FS><<<grid, kSitesPerPass * C, ring_bytes<C>(FS), e->stream>>>(a);
      }
      else
      {
        if (a.want_lnl) newton_sweep_kernel<C, true, false><<<grid, kSitesPerPass * C, sweep_ring_bytes<C>(), e->stream>>>(a);
        else newton_sweep_kernel<C, false, false><<<grid, kSitesPerPass * C, sweep_ring_bytes<C>(), e->stream>>>(a);
      }
    }
    { LaunchScope ls(e, PTGPU_KERNEL_NEWTON); newton_reduce_kernel<false><<<a.B, 256, 0, e->stream>>>(a, e->d_red); }
    CU(cudaGetLastError());
    if ((rc = shard_sum(e, e->d_red, (size_t)a.B * 3))) return rc;
    CU(cudaMemsetAsync(e->d_active, 0, 2 * sizeof(unsigned int), e->stream));
    { LaunchScope ls(e, PTGPU_KERNEL_NEWTON); newton_update_kernel<<<(a.B + 127) / 128, 128, 0, e->stream>>>(a, e->d_red, e->d_active); }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync((void *)active_h, e->d_active, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    pending_lnl = active_h[1];
    if (active_h[0] == 0) break;
  }
  if (!a.want_lnl) return PTGPU_OK;
  // every shard sees the same flags (same sums, same updates), so all of them
  // take the same branch here
  if (!pending_lnl) return PTGPU_OK;
  {
    LaunchScope ls(e, PTGPU_KERNEL_NEWTON);
    if (deep) newton_sweep_kernel<C, true, true, FS><<<grid, kSitesPerPass * C, ring_bytes<C>(FS), e->stream>>>(a);
    else newton_sweep_kernel<C, true, true><<<grid, kSitesPerPass * C, sweep_ring_bytes<C>(), e->stream>>>(a);
  }
  { LaunchScope ls(e, PTGPU_KERNEL_NEWTON); newton_reduce_kernel<true><<<a.B, 256, 0, e->stream>>>(a, e->d_red); }
  CU(cudaGetLastError());
  if ((rc = shard_sum(e, e->d_red, (size_t)a.B * 3))) return rc;
  { LaunchScope ls(e, PTGPU_KERNEL_NEWTON); newton_store_lnl_kernel<<<(a.B + 127) / 128, 128, 0, e->stream>>>(a, e->d_red); }
  CU(cudaGetLastError());
  return PTGPU_OK;
}

int run_newton(ptgpu_engine *e, NewtonArgs &a)
{
  a.vec_end = e->vec_end;
  a.lnl_margin = e->lnl_margin;
  if (is_sharded(e)) { DISPATCH_CATS(e, return newton_split<C>(e, a)); }
  DISPATCH_CATS(e, return launch_newton<C>(e, a));
  return PTGPU_OK;
}

} // namespace

extern "C" {

int ptgpu_update_prob_matrices(ptgpu_engine *e, const unsigned int *matrix_indices,
                               const double *branch_lengths, unsigned int count)
{
  if (!e || (count && (!matrix_indices || !branch_lengths)))
    return fail(PTGPU_ERR_ARG, "ptgpu_update_prob_matrices: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  if (!count) return PTGPU_OK;
  for (unsigned int i = 0; i < count; ++i)
  {
    int rc = check_matrix(e, matrix_indices[i], "update_prob_matrices");
    if (rc) return rc;
    if (!(branch_lengths[i] >= 0))
      return fail(PTGPU_ERR_ARG, "negative or NaN branch length %g", branch_lengths[i]);
  }
  CU(cudaSetDevice(e->device));
  int rc;
  if ((rc = ensure(&e->d_midx, &e->midx_cap, count))) return rc;
  if ((rc = ensure(&e->d_mlen, &e->mlen_cap, count))) return rc;
  if ((rc = upload(e, e->d_midx, matrix_indices, count * sizeof(unsigned int)))) return rc;
  if ((rc = upload(e, e->d_mlen, branch_lengths, count * sizeof(double)))) return rc;
  DISPATCH_CATS(e, rc = launch_pmatrix<C>(e, e->d_midx, e->d_mlen, count));
  return rc;
}

int ptgpu_update_partials(ptgpu_engine *e, const ptgpu_operation_t *ops, unsigned int count)
{
  if (!e || (count && !ops)) return fail(PTGPU_ERR_ARG, "ptgpu_update_partials: null argument");
  if (!count) return PTGPU_OK;
  int rc;
  for (unsigned int i = 0; i < count; ++i)
    if ((rc = check_op(e, ops[i], true))) return rc;
  CU(cudaSetDevice(e->device));
  if ((rc = ensure(&e->d_ops, &e->ops_cap, count))) return rc;
  if ((rc = upload(e, e->d_ops, ops, count * sizeof(DevOp)))) return rc;
  {
    LaunchScope ls(e, PTGPU_KERNEL_PARTIALS);
    if (e->partials_version == 3)
    { DISPATCH_CATS(e, (partials_kernel<C><<<e->nheavy, kHeavySites * C, 0, e->stream>>>(e->dev, e->d_ops, count))); }
    else if (e->nheavy < 4u * (unsigned int)e->sm_count)
    {
      // too few 128-site tiles to fill the device: 64-site tiles, twice the warps (bit-identical)
      const unsigned int tiles = (e->sites + 2 * kHeavySites - 1) / (2 * kHeavySites);
      DISPATCH_CATS(e, (partials_kernel_v4<C, 2><<<tiles, kHeavySites * C, 0, e->stream>>>(e->dev, e->d_ops, count)));
    }
    else
    { DISPATCH_CATS(e, (partials_kernel_v4<C><<<e->nheavy, kHeavySites * C, 0, e->stream>>>(e->dev, e->d_ops, count))); }
  }
  CU(cudaGetLastError());
  e->counters[1] += (unsigned long long)count * e->sites;
  return PTGPU_OK;
}

int ptgpu_compute_edge_loglikelihood(ptgpu_engine *e, unsigned int parent_clv_index,
                                     int parent_scaler_index, unsigned int child_clv_index,
                                     int child_scaler_index, unsigned int matrix_index,
                                     double *persite_lnl, double *lnl_out)
{
  if (!e || !lnl_out) return fail(PTGPU_ERR_ARG, "ptgpu_compute_edge_loglikelihood: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  int rc;
  if ((rc = check_clv(e, parent_clv_index, "edge_lnl.parent"))) return rc;
  if ((rc = check_clv(e, child_clv_index, "edge_lnl.child"))) return rc;
  if ((rc = check_scaler(e, parent_scaler_index, "edge_lnl.parent"))) return rc;
  if ((rc = check_scaler(e, child_scaler_index, "edge_lnl.child"))) return rc;
  if ((rc = check_matrix(e, matrix_index, "edge_lnl"))) return rc;
  CU(cudaSetDevice(e->device));
  double *d_persite = nullptr;
  if (persite_lnl)
  {
    if (!e->d_persite) CU(PT_MALLOC(e, &e->d_persite, (size_t)e->sites * sizeof(double)));
    d_persite = e->d_persite;
  }
  double *result = (double *)(e->stg.dev + kStgResult);
  if (is_sharded(e))
  {
    if ((rc = ensure(&e->d_red, &e->red_cap, 2))) return rc;
    result = e->d_red;  // shard partial stays on the device for the hook
  }
  {
    LaunchScope ls(e, PTGPU_KERNEL_EDGE_LNL);
    DISPATCH_CATS(e, (edge_lnl_kernel<C><<<e->ntiles, kSitesPerPass * C, 0, e->stream>>>(
                         e->dev, e->model, parent_clv_index, parent_scaler_index, child_clv_index,
                         child_scaler_index, matrix_index, d_persite, e->d_partials, e->d_counter, result)));
  }
  CU(cudaGetLastError());
  if (persite_lnl)
    CU(cudaMemcpyAsync(persite_lnl, d_persite, (size_t)e->sites * sizeof(double),
                       cudaMemcpyDeviceToHost, e->stream));
  if (is_sharded(e))
  {
    if ((rc = shard_sum(e, e->d_red, 1))) return rc;
    CU(cudaMemcpyAsync(e->stg.host + kStgResult, e->d_red, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  }
  CU(cudaStreamSynchronize(e->stream));
  *lnl_out = *(volatile double *)(e->stg.host + kStgResult);
  return shard_check(e);
}

int ptgpu_update_sumtable(ptgpu_engine *e, unsigned int parent_clv_index,
                          unsigned int child_clv_index, int parent_scaler_index,
                          int child_scaler_index)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_update_sumtable: null engine");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  int rc;
  if ((rc = check_clv(e, parent_clv_index, "sumtable.parent"))) return rc;
  if ((rc = check_clv(e, child_clv_index, "sumtable.child"))) return rc;
  (void)parent_scaler_index; (void)child_scaler_index; // per-site scalers do not enter the sumtable
  CU(cudaSetDevice(e->device));
  {
    LaunchScope ls(e, PTGPU_KERNEL_SUMTABLE);
    DISPATCH_CATS(e, (sumtable_kernel<C><<<e->ntiles, kSitesPerPass * C, 0, e->stream>>>(
                         e->dev, e->model, parent_clv_index, child_clv_index, e->d_sumtable)));
  }
  CU(cudaGetLastError());
  return PTGPU_OK;
}

static int newton_single(ptgpu_engine *e, double len0, unsigned int max_iter, double eps,
                         double lambda, NewtonState *out, int mode = 0, double xmin = 0.0,
                         double xmax = 0.0)
{
  NewtonState init{};
  init.len = len0;
  int rc;
  if ((rc = upload(e, e->d_state, &init, sizeof init))) return rc;
  NewtonArgs a{};
  a.m = e->model;
  a.sumtables = e->d_sumtable;
  a.entry_stride = 0;
  a.sclr = nullptr;
  a.st = e->d_state;
  a.partials = e->d_partials;
  a.sites = e->sites; a.B = 1; a.ntiles = e->ntiles; a.chunk_tiles = e->chunk_tiles; a.nchunks = e->nchunks; a.max_iter = max_iter;
  a.eps = eps; a.lambda = lambda; a.want_lnl = 0;
  a.mode = mode; a.xmin = xmin; a.xmax = xmax;
  if ((rc = run_newton(e, a))) return rc;
  e->counters[2] += max_iter; // upper bound; exact count is returned in state.iters
  (void)out;
  return PTGPU_OK;
}

int ptgpu_compute_likelihood_derivatives(ptgpu_engine *e, double branch_length, double *d_f,
                                         double *dd_f)
{
  if (!e || !d_f || !dd_f) return fail(PTGPU_ERR_ARG, "ptgpu_compute_likelihood_derivatives: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  CU(cudaSetDevice(e->device));
  int rc;
  if ((rc = newton_single(e, branch_length, 1, 0.0, 0.0, nullptr))) return rc;
  NewtonState *hs = (NewtonState *)(e->stg.host + kStgNewton);
  CU(cudaMemcpyAsync(hs, e->d_state, sizeof(NewtonState), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  *d_f = hs->d1;
  *dd_f = hs->d2;
  return shard_check(e);
}

int ptgpu_optimize_branch(ptgpu_engine *e, unsigned int parent_clv_index,
                          int parent_scaler_index, unsigned int child_clv_index,
                          int child_scaler_index, unsigned int matrix_index, double length0,
                          unsigned int max_iter, double epsilon, double map_lambda,
                          double *length_out, int *iterations)
{
  if (!e || !length_out) return fail(PTGPU_ERR_ARG, "ptgpu_optimize_branch: null argument");
  int rc;
  if ((rc = check_matrix(e, matrix_index, "optimize_branch"))) return rc;
  if ((rc = ptgpu_update_sumtable(e, parent_clv_index, child_clv_index, parent_scaler_index,
                                  child_scaler_index)))
    return rc;
  if ((rc = newton_single(e, length0, max_iter, epsilon, map_lambda, nullptr))) return rc;
  // UpdateBranchLength (PTPLL/pll_partition.cpp:395-406): the edge's P matrix at
  // the new length, straight from the device-resident state
  *(unsigned int *)(e->stg.host + kStgMatrixIdx) = matrix_index;
  CU(cudaMemcpyAsync(e->d_idx1, e->stg.host + kStgMatrixIdx, sizeof(unsigned int), cudaMemcpyHostToDevice, e->stream));
  DISPATCH_CATS(e, rc = launch_pmatrix<C>(e, e->d_idx1, &e->d_state->len, 1));
  if (rc) return rc;
  NewtonState *hs = (NewtonState *)(e->stg.host + kStgNewton);
  CU(cudaMemcpyAsync(hs, e->d_state, sizeof(NewtonState), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  *length_out = hs->len;
  if (iterations) *iterations = hs->iters;
  if ((rc = shard_check(e))) return rc;
  if (hs->status == 1) return fail(PTGPU_ERR_NOT_FINITE, "d1 is not finite");
  if (hs->status == 2) return fail(PTGPU_ERR_NOT_FINITE, "d2 is not finite");
  return PTGPU_OK;
}

int ptgpu_optimize_branch_bounded(ptgpu_engine *e, unsigned int parent_clv_index,
                                  int parent_scaler_index, unsigned int child_clv_index,
                                  int child_scaler_index, double xguess, double xmin,
                                  double xmax, double tolerance, unsigned int max_iter,
                                  double *length_out, int *status, int *iterations)
{
  if (!e || !length_out || !status)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_branch_bounded: null argument");
  if (!(xmin <= xmax) || !(tolerance > 0) || !max_iter)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_branch_bounded: bad bounds / tolerance");
  int rc;
  if ((rc = ptgpu_update_sumtable(e, parent_clv_index, child_clv_index, parent_scaler_index,
                                  child_scaler_index)))
    return rc;
  // rts = xguess clamped into [xmin, xmax] (PLLMOD/optimize/opt_algorithms.c:88-92)
  double rts = xguess;
  if (rts < xmin) rts = xmin;
  if (rts > xmax) rts = xmax;
  if ((rc = newton_single(e, rts, max_iter, tolerance, 0.0, nullptr, 1, xmin, xmax))) return rc;
  NewtonState *hs = (NewtonState *)(e->stg.host + kStgNewton);
  CU(cudaMemcpyAsync(hs, e->d_state, sizeof(NewtonState), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  *length_out = hs->len;
  *status = hs->status;
  if (iterations) *iterations = hs->iters;
  return shard_check(e);
}

int ptgpu_score_nni_batch(ptgpu_engine *e, const ptgpu_nni_candidate_t *cands,
                          unsigned int count, unsigned int max_iter, double epsilon,
                          double map_lambda, double *length_out, double *lnl_out,
                          int *status_out)
{
  if (!e || (count && (!cands || !length_out || !lnl_out)))
    return fail(PTGPU_ERR_ARG, "ptgpu_score_nni_batch: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  if (!count) return PTGPU_OK;
  int rc;
  for (unsigned int i = 0; i < count; ++i)
  {
    if ((rc = check_op(e, cands[i].p, false))) return rc;
    if ((rc = check_op(e, cands[i].q, false))) return rc;
    if (!(cands[i].length0 >= 0)) return fail(PTGPU_ERR_ARG, "candidate %u: bad start length", i);
  }
  CU(cudaSetDevice(e->device));
  // compact sumtables: tile-blocked, 3 C + 1 doubles per site (never more than the full form's
  // 4 C except on alignments shorter than a tile)
  const int compact = (e->compact_sumtable && e->zero_col >= 0 && e->prepare_version != 1) ? e->zero_col + 1 : 0;
  const size_t stride = std::max(e->dev.clv_stride,
                                 compact ? (size_t)e->ntiles * kTileSites * (3 * e->cats + 1) : (size_t)0);
  // candidates are processed in chunks whose sumtables fit the scratch budget
  // chunk partials: one per chunk, or one per chunk and WARP for the site-per-thread sweep
  const size_t nparts = std::max<size_t>(e->ntiles, (size_t)e->nchunks * 16);
  const size_t per_cand = stride * sizeof(double) + (size_t)e->sites * sizeof(unsigned int) +
                          nparts * 3 * sizeof(double) + sizeof(NewtonState) + sizeof(DevGroup);
  size_t chunk = count;
  if (chunk > e->bcap)
  {
    // the scratch has to grow: as many candidates per launch as 80 % of the free memory holds
    // (asked only now -- cudaMemGetInfo is a heavyweight driver call, not one for every batch)
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    const size_t have = e->bcap * per_cand;
    const size_t budget = std::max(have, (size_t)((free_b + have) * 0.8));
    chunk = std::min<size_t>(count, std::max<size_t>(1, budget / per_cand));
    // all shards must cut a batch into the same launches; free memory may differ between them
    if (is_sharded(e) && chunk < count)
      return fail(PTGPU_ERR_STATE, "site shards: a batch of %u candidates does not fit this shard's memory", count);
  }
  if (chunk > e->bcap)
  {
    PT_FREE(e, e->d_bsum); PT_FREE(e, e->d_bsclr); PT_FREE(e, e->d_bstate); PT_FREE(e, e->d_bpartials); PT_FREE(e, e->d_bcands);
    e->d_bsum = nullptr; e->d_bsclr = nullptr; e->d_bstate = nullptr; e->d_bpartials = nullptr; e->d_bcands = nullptr;
    e->bcap = 0;
    CU(PT_MALLOC(e, &e->d_bsum, chunk * stride * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_bsclr, chunk * (size_t)e->sites * sizeof(unsigned int)));
    CU(PT_MALLOC(e, &e->d_bstate, chunk * sizeof(NewtonState)));
    CU(PT_MALLOC(e, &e->d_bpartials, chunk * nparts * 3 * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_bcands, chunk * sizeof(DevGroup)));
    e->bcap = chunk;
  }
  // per-candidate Newton states go up and come back through the pinned arena
  if ((rc = pin_reserve(e, chunk * sizeof(NewtonState)))) return rc;
  NewtonState *hst = reinterpret_cast<NewtonState *>(e->h_pin);
  for (unsigned int base = 0; base < count; base += (unsigned int)chunk)
  {
    const unsigned int nb = (unsigned int)std::min<size_t>(chunk, count - base);
    for (unsigned int i = 0; i < nb; ++i)
    {
      hst[i] = NewtonState{};
      hst[i].len = cands[base + i].length0;
    }
    // candidates that read the same four neighbours (the LEFT and RIGHT move
    // across one edge) share one group: their matrix-vector products are done once
    std::vector<DevGroup> groups;
    groups.reserve(nb);
    auto child = [](const ptgpu_operation_t &op, int which) {
      DevChild c;
      c.clv = which ? op.child2_clv_index : op.child1_clv_index;
      c.scaler = which ? op.child2_scaler_index : op.child1_scaler_index;
      c.mat = which ? op.child2_matrix_index : op.child1_matrix_index;
      c.pad = 0;
      return c;
    };
    auto same = [](const DevChild &a, const DevChild &b) {
      return a.clv == b.clv && a.scaler == b.scaler && a.mat == b.mat;
    };
    for (unsigned int i = 0; i < nb; ++i)
    {
      const ptgpu_nni_candidate_t &c0 = cands[base + i];
      DevGroup g{};
      g.ch[0] = child(c0.p, 0); g.ch[1] = child(c0.p, 1);
      g.ch[2] = child(c0.q, 0); g.ch[3] = child(c0.q, 1);
      g.ncand = 1;
      g.out[0] = i;
      if (i + 1 < nb)
      {
        // second candidate must combine (ch3, ch1) on one side and (ch2, ch0) on the other;
        // child order within a side and which side is "p" for the swap partner are free
        // up to relabelling ch0<->ch1 / ch2<->ch3 of the first candidate
        const ptgpu_nni_candidate_t &c1 = cands[base + i + 1];
        const DevChild a0 = child(c1.p, 0), a1 = child(c1.p, 1), b0 = child(c1.q, 0), b1 = child(c1.q, 1);
        for (int flip = 0; flip < 4 && g.ncand == 1; ++flip)
        {
          DevGroup t = g;
          if (flip & 1) std::swap(t.ch[0], t.ch[1]);
          if (flip & 2) std::swap(t.ch[2], t.ch[3]);
          const bool pside = (same(a0, t.ch[3]) && same(a1, t.ch[1])) || (same(a0, t.ch[1]) && same(a1, t.ch[3]));
          const bool qside = (same(b0, t.ch[2]) && same(b1, t.ch[0])) || (same(b0, t.ch[0]) && same(b1, t.ch[2]));
          if (pside && qside)
          {
            g = t;
            g.ncand = 2;
            g.out[1] = i + 1;
          }
        }
      }
      groups.push_back(g);
      if (g.ncand == 2) ++i;
    }
    const unsigned int ng = (unsigned int)groups.size();
    if ((rc = upload(e, e->d_bcands, groups.data(), ng * sizeof(DevGroup)))) return rc;
    CU(cudaMemcpyAsync(e->d_bstate, hst, nb * sizeof(NewtonState), cudaMemcpyHostToDevice, e->stream));
    {
      LaunchScope ls(e, PTGPU_KERNEL_NNI_PREPARE);
      if (e->prepare_version == 1)
      { DISPATCH_CATS(e, (nni_prepare_kernel<C><<<e->nheavy * ng, kHeavySites * C, sizeof(PrepSmem<C>), e->stream>>>(
                           e->dev, e->model, e->d_bcands, ng, e->d_bsum, stride, e->d_bsclr))); }
      else
      { DISPATCH_CATS(e, (nni_prepare_kernel_v2<C><<<e->nheavy * ng, kHeavySites * C, sizeof(PrepSmem2<C>), e->stream>>>(
                           e->dev, e->model, e->d_bcands, ng, e->d_bsum, stride, e->d_bsclr, compact))); }
    }
    CU(cudaGetLastError());
    NewtonArgs a{};
    a.compact = compact;
    a.m = e->model;
    a.sumtables = e->d_bsum;
    a.entry_stride = stride;
    a.sclr = e->d_bsclr;
    a.st = e->d_bstate;
    a.partials = e->d_bpartials;
    a.sites = e->sites; a.B = nb; a.ntiles = e->ntiles; a.chunk_tiles = e->chunk_tiles; a.nchunks = e->nchunks; a.max_iter = max_iter;
    a.eps = epsilon; a.lambda = map_lambda; a.want_lnl = 1;
    if ((rc = run_newton(e, a))) return rc;
    CU(cudaMemcpyAsync(hst, e->d_bstate, nb * sizeof(NewtonState), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    if ((rc = shard_check(e))) return rc;
    for (unsigned int i = 0; i < nb; ++i)
    {
      length_out[base + i] = hst[i].len;
      lnl_out[base + i] = hst[i].lnl;
      if (status_out) status_out[base + i] = hst[i].status ? PTGPU_ERR_NOT_FINITE : PTGPU_OK;
      e->counters[2] += (unsigned long long)hst[i].iters;
      e->lnl_sweeps += (unsigned long long)hst[i].lnl_sweeps;
      e->final_sweeps += (unsigned long long)hst[i].final_sweeps;
    }
    if (getenv("PTGPU_DEBUG"))
      fprintf(stderr, "[ptgpu] batch of %u: %llu lnL-carrying sweeps, %llu extra final sweeps so far\n", nb,
              e->lnl_sweeps, e->final_sweeps);
    e->counters[3] += nb;
  }
  return PTGPU_OK;
}

int ptgpu_set_allreduce(ptgpu_engine *e, ptgpu_allreduce_fn fn, void *ctx)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_set_allreduce: null engine");
  e->allreduce = fn;
  e->allreduce_ctx = ctx;
  return PTGPU_OK;
}

int ptgpu_set_shard(ptgpu_engine *e, unsigned long long site_offset, unsigned long long global_sites)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_set_shard: null engine");
  if (site_offset + e->sites > global_sites)
    return fail(PTGPU_ERR_ARG, "ptgpu_set_shard: shard [%llu, %llu) exceeds %llu sites", site_offset,
                site_offset + e->sites, global_sites);
  // the reference's derivative loop treats the last global_sites % 4 sites
  // differently (scalar tail, LIBPLL/core_derivatives_avx.c:1148-1156)
  const unsigned long long gfull = global_sites / 4 * 4;
  e->vec_end = gfull <= site_offset ? 0ull : gfull - site_offset;
  return PTGPU_OK;
}

int ptgpu_get_clv(ptgpu_engine *e, unsigned int clv_index, double *out)
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_get_clv: null argument");
  int rc = check_clv(e, clv_index, "get_clv");
  if (rc) return rc;
  CU(cudaSetDevice(e->device));
  if (clv_index < e->tips)
  {
    std::vector<uint8_t> m(e->sites);
    CU(cudaMemcpyAsync(m.data(), e->d_tips + (size_t)clv_index * e->dev.tip_stride, e->sites, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    for (unsigned int s = 0; s < e->sites; ++s)
      for (unsigned int c = 0; c < e->cats; ++c)
        for (int j = 0; j < 4; ++j) out[((size_t)s * e->cats + c) * 4 + j] = (m[s] >> j) & 1 ? 1.0 : 0.0;
    return PTGPU_OK;
  }
  CU(cudaMemcpyAsync(out, e->dev.clv + (size_t)(clv_index - e->tips) * e->dev.clv_stride,
                     e->dev.clv_stride * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

int ptgpu_set_clv(ptgpu_engine *e, unsigned int clv_index, const double *in)
{
  if (!e || !in) return fail(PTGPU_ERR_ARG, "ptgpu_set_clv: null argument");
  int rc = check_clv(e, clv_index, "set_clv");
  if (rc) return rc;
  if (clv_index < e->tips) return fail(PTGPU_ERR_ARG, "set_clv: %u is a tip", clv_index);
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(e->dev.clv + (size_t)(clv_index - e->tips) * e->dev.clv_stride, in,
                     e->dev.clv_stride * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

int ptgpu_get_scaler(ptgpu_engine *e, int scaler_index, unsigned int *out)
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_get_scaler: null argument");
  if (scaler_index < 0 || (unsigned int)scaler_index >= e->scale_buffers)
    return fail(PTGPU_ERR_ARG, "get_scaler: index %d out of range", scaler_index);
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(out, e->dev.scalers + (size_t)scaler_index * e->dev.scaler_stride,
                     (size_t)e->sites * sizeof(unsigned int), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

int ptgpu_get_pmatrix(ptgpu_engine *e, unsigned int matrix_index, double *out)
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_get_pmatrix: null argument");
  int rc = check_matrix(e, matrix_index, "get_pmatrix");
  if (rc) return rc;
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(out, e->dev.pmat + (size_t)matrix_index * e->cats * 16,
                     (size_t)e->cats * 16 * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

int ptgpu_set_pmatrix(ptgpu_engine *e, unsigned int matrix_index, const double *in)
{
  if (!e || !in) return fail(PTGPU_ERR_ARG, "ptgpu_set_pmatrix: null argument");
  int rc = check_matrix(e, matrix_index, "set_pmatrix");
  if (rc) return rc;
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(e->dev.pmat + (size_t)matrix_index * e->cats * 16, in,
                     (size_t)e->cats * 16 * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

int ptgpu_get_sumtable(ptgpu_engine *e, double *out)
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_get_sumtable: null argument");
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(out, e->d_sumtable, e->dev.clv_stride * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

unsigned int ptgpu_batch_sumtable_doubles(ptgpu_engine *e)
{
  if (!e) return 0;
  const bool compact = e->compact_sumtable && e->zero_col >= 0 && e->prepare_version != 1;
  return compact ? 3 * e->cats + 1 : 4 * e->cats;
}

int ptgpu_get_counters(ptgpu_engine *e, unsigned long long out[6])
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_get_counters: null argument");
  for (int i = 0; i < 4; ++i) out[i] = e->counters[i];
  out[4] = e->lnl_sweeps;
  out[5] = e->final_sweeps;
  return PTGPU_OK;
}

int ptgpu_set_profiling(ptgpu_engine *e, int on)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_set_profiling: null engine");
  e->profiling = on != 0;
  return PTGPU_OK;
}

int ptgpu_get_kernel_times(ptgpu_engine *e, double ms_out[PTGPU_KERNEL_KINDS],
                           unsigned long long launches_out[PTGPU_KERNEL_KINDS], int reset)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_get_kernel_times: null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  for (auto &p : e->pending)
  {
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, p.second.first, p.second.second));
    e->kernel_ms[p.first] += ms;
    e->event_pool.push_back(p.second.first);
    e->event_pool.push_back(p.second.second);
  }
  e->pending.clear();
  for (int k = 0; k < PTGPU_KERNEL_KINDS; ++k)
  {
    if (ms_out) ms_out[k] = e->kernel_ms[k];
    if (launches_out) launches_out[k] = e->kernel_launches[k];
    if (reset) { e->kernel_ms[k] = 0; e->kernel_launches[k] = 0; }
  }
  return PTGPU_OK;
}

void *ptgpu_stream(ptgpu_engine *e) { return e ? (void *)e->stream : nullptr; }

int ptgpu_synchronize(ptgpu_engine *e)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_synchronize: null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  return PTGPU_OK;
}

} // extern "C"


// ---- the engines' own exchange between site shards (kernels.cuh, ShardComm) --------------
namespace {

int blo_prepare_kernel(ptgpu_engine *e);

// With lazy module loading (the CUDA 12 default) the FIRST launch of a kernel loads it, which
// can synchronise the context -- fatal if a peer shard's kernel on the same device is already
// waiting for this one.  Every kernel a shard can run is therefore loaded before the shards
// are connected.
template <int C>
int shard_preload()
{
  cudaFuncAttributes fa;
  constexpr int FS = few_work_stages<C>();
#define PT_LOAD(fn) CU(cudaFuncGetAttributes(&fa, fn))
  PT_LOAD(tip_encode_kernel); PT_LOAD(tip_encode_batch_kernel); PT_LOAD(shard_allreduce_kernel);
  PT_LOAD(sum_partials_kernel); PT_LOAD(newton_update_kernel); PT_LOAD(newton_store_lnl_kernel);
  PT_LOAD(newton_reduce_kernel<false>); PT_LOAD(newton_reduce_kernel<true>);
  PT_LOAD(pmatrix_kernel<C>); PT_LOAD(partials_kernel<C>); PT_LOAD(partials_kernel_v4<C>); PT_LOAD((partials_kernel_v4<C, 2>));
  PT_LOAD(edge_lnl_kernel<C>); PT_LOAD(sumtable_kernel<C>);
  PT_LOAD(nni_prepare_kernel<C>); PT_LOAD(nni_prepare_kernel_v2<C>);
  PT_LOAD((newton_kernel<C, false>)); PT_LOAD((newton_kernel<C, true>));
  PT_LOAD((newton_kernel<C, false, FS>)); PT_LOAD((newton_kernel<C, true, FS>));
  PT_LOAD((newton_sweep_kernel<C, false, false>)); PT_LOAD((newton_sweep_kernel<C, true, false>));
  PT_LOAD((newton_sweep_kernel<C, true, true>));
  PT_LOAD((newton_sweep_kernel<C, false, false, FS>)); PT_LOAD((newton_sweep_kernel<C, true, false, FS>));
  PT_LOAD((newton_sweep_kernel<C, true, true, FS>));
  PT_LOAD((blo_kernel<C, false>)); PT_LOAD((blo_kernel<C, true>));
#undef PT_LOAD
  return PTGPU_OK;
}

int shard_alloc(ptgpu_engine *e)
{
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  if (!e->d_mailbox) CU(cudaMalloc(&e->d_mailbox, shard_mailbox_bytes()));
  if (!e->d_seq) CU(cudaMalloc(&e->d_seq, sizeof(unsigned long long)));
  CU(cudaMemset(e->d_mailbox, 0, shard_mailbox_bytes()));
  CU(cudaMemset(e->d_seq, 0, sizeof(unsigned long long)));
  *(volatile int *)(e->stg.host + kStgShardErr) = 0;
  int rc = PTGPU_OK;
  DISPATCH_CATS(e, rc = shard_preload<C>());
  if (rc) return rc;
  if ((rc = blo_prepare_kernel(e))) return rc;
  if (!e->d_active) CU(PT_MALLOC(e, &e->d_active, 2 * sizeof(unsigned int)));
  CU(cudaDeviceSynchronize());
  return PTGPU_OK;
}

void shard_fill(ptgpu_engine *e, unsigned int rank, unsigned int nranks)
{
  e->comm.rank = (int)rank;
  e->comm.nranks = (int)nranks;
  e->comm.seq = e->d_seq;
  e->comm.error = (int *)(e->stg.dev + kStgShardErr);
  unsigned long long ms = 20000;
  if (const char *v = getenv("PTGPU_SHARD_TIMEOUT_MS")) ms = std::max(1ull, strtoull(v, nullptr, 10));
  e->comm.timeout_ns = ms * 1000000ull;
}

void shard_close(ptgpu_engine *e)
{
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (void *&p : e->ipc_open)
    if (p) { cudaIpcCloseMemHandle(p); p = nullptr; }
  e->comm = ShardComm{};
  e->same_device_shards = 1;
}

} // namespace

extern "C" {

int ptgpu_shard_export(ptgpu_engine *e, ptgpu_shard_handle_t *out)
{
  if (!e || !out) return fail(PTGPU_ERR_ARG, "ptgpu_shard_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(ptgpu_shard_handle_t), "handle size");
  shard_close(e);
  int rc;
  if ((rc = shard_alloc(e))) return rc;
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, e->d_mailbox));
  memset(out, 0, sizeof *out);
  memcpy(out->bytes, &h, sizeof h);
  return PTGPU_OK;
}

int ptgpu_shard_connect(ptgpu_engine *e, unsigned int rank, unsigned int nranks,
                        const ptgpu_shard_handle_t *handles)
{
  if (!e || !handles) return fail(PTGPU_ERR_ARG, "ptgpu_shard_connect: null argument");
  if (nranks < 1 || nranks > (unsigned int)kMaxShards || rank >= nranks)
    return fail(PTGPU_ERR_ARG, "ptgpu_shard_connect: rank %u of %u (at most %d shards)", rank, nranks, kMaxShards);
  if (!e->d_mailbox) return fail(PTGPU_ERR_STATE, "ptgpu_shard_connect: call ptgpu_shard_export first");
  if (e->allreduce) return fail(PTGPU_ERR_STATE, "ptgpu_shard_connect: an allreduce hook is set");
  CU(cudaSetDevice(e->device));
  for (unsigned int r = 0; r < nranks; ++r)
  {
    if (r == rank) { e->comm.box[r] = e->d_mailbox; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handles[r].bytes, sizeof h);
    void *p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    e->ipc_open[r] = p;
    e->comm.box[r] = (unsigned long long *)p;
  }
  shard_fill(e, rank, nranks);
  return PTGPU_OK;
}

int ptgpu_comm_init(ptgpu_engine *const *shards, unsigned int nshards)
{
  if (!shards || nshards < 1 || nshards > (unsigned int)kMaxShards)
    return fail(PTGPU_ERR_ARG, "ptgpu_comm_init: 1 to %d shards", kMaxShards);
  int rc;
  for (unsigned int r = 0; r < nshards; ++r)
  {
    if (!shards[r]) return fail(PTGPU_ERR_ARG, "ptgpu_comm_init: shard %u is null", r);
    if (shards[r]->allreduce) return fail(PTGPU_ERR_STATE, "ptgpu_comm_init: shard %u has an allreduce hook", r);
    for (unsigned int k = 0; k < r; ++k)
      if (shards[k] == shards[r]) return fail(PTGPU_ERR_ARG, "ptgpu_comm_init: shard %u listed twice", r);
    shard_close(shards[r]);
    if ((rc = shard_alloc(shards[r]))) return rc;
  }
  // every device maps every other one (NVLink / NVSwitch peer access)
  for (unsigned int a = 0; a < nshards; ++a)
    for (unsigned int b = 0; b < nshards; ++b)
    {
      const int da = shards[a]->device, db = shards[b]->device;
      if (da == db) continue;
      int can = 0;
      CU(cudaDeviceCanAccessPeer(&can, da, db));
      if (!can) return fail(PTGPU_ERR_CUDA, "ptgpu_comm_init: device %d cannot access device %d", da, db);
      CU(cudaSetDevice(da));
      const cudaError_t err = cudaDeviceEnablePeerAccess(db, 0);
      if (err == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (err != cudaSuccess)
        return fail(PTGPU_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", da, db, cudaGetErrorString(err));
    }
  for (unsigned int a = 0; a < nshards; ++a)
  {
    int same = 0;
    for (unsigned int b = 0; b < nshards; ++b)
    {
      shards[a]->comm.box[b] = shards[b]->d_mailbox;
      same += shards[b]->device == shards[a]->device;
    }
    shards[a]->same_device_shards = same;
    shard_fill(shards[a], a, nshards);
  }
  return PTGPU_OK;
}

int ptgpu_shard_disconnect(ptgpu_engine *e)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_shard_disconnect: null engine");
  shard_close(e);
  return PTGPU_OK;
}

int ptgpu_shard_status(ptgpu_engine *e, int *nranks_out, int *error_out, unsigned long long *exchanges_out)
{
  if (!e) return fail(PTGPU_ERR_ARG, "ptgpu_shard_status: null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  if (nranks_out) *nranks_out = e->comm.nranks;
  if (error_out) *error_out = *(volatile int *)(e->stg.host + kStgShardErr);
  if (exchanges_out)
  {
    *exchanges_out = 0;
    if (e->d_seq) CU(cudaMemcpy(exchanges_out, e->d_seq, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  }
  return PTGPU_OK;
}

} // extern "C"

namespace {

constexpr unsigned int kBloCluster = 8;   // CTAs of the single-cluster variant (portable maximum)

template <int C>
int blo_occupancy(int *out)
{
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(out, blo_kernel<C, false>, kSitesPerPass * C,
                                                   sweep_ring_bytes<C>()));
  return PTGPU_OK;
}

template <int C>
int launch_blo(ptgpu_engine *e, BloArgs &a, unsigned int grid)
{
  void *params[] = {&a};
  LaunchScope ls(e, PTGPU_KERNEL_NEWTON);
  CU(cudaLaunchCooperativeKernel((void *)blo_kernel<C, false>, dim3(grid), dim3(kSitesPerPass * C), params,
                                 sweep_ring_bytes<C>(), e->stream));
  return PTGPU_OK;
}

// the whole grid as one thread-block cluster; barriers are the hardware cluster barrier
template <int C>
int launch_blo_cluster(ptgpu_engine *e, BloArgs &a)
{
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(kBloCluster);
  cfg.blockDim = dim3(kSitesPerPass * C);
  cfg.dynamicSmemBytes = sweep_ring_bytes<C>();
  cfg.stream = e->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kBloCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  LaunchScope ls(e, PTGPU_KERNEL_NEWTON);
  CU(cudaLaunchKernelEx(&cfg, blo_kernel<C, true>, a));
  return PTGPU_OK;
}

// one request to the persistent step-program kernel (blo.cuh)
struct BloRequest
{
  unsigned int count, n_private_clvs, n_lengths, n_prologue, n_body, n_bodies;
  const int *step_types;
  const ptgpu_blo_step_t *steps;
  const ptgpu_blo_edge_t *smooth_edges, *final_edges;
  const double *lengths0;
  int procedure;             // BloArgs::procedure
  double xmin, xmax, xtol, tolerance, lambda;
  unsigned int smoothings, max_newton;
  const unsigned int *prior_slots;
  unsigned int n_prior;
  double *lengths_out, *lnl_out;
  int *status_out, *smoothings_out;
};

int blo_prepare_kernel(ptgpu_engine *e)
{
  if (e->blo_blocks_per_sm) return PTGPU_OK;
  int occ = 0, rc = PTGPU_OK;
  cudaError_t ea = cudaSuccess;
  switch (e->cats)
  {
    case 1: ea = cudaFuncSetAttribute(blo_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_ring_bytes<1>()); rc = blo_occupancy<1>(&occ); break;
    case 2: ea = cudaFuncSetAttribute(blo_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_ring_bytes<2>()); rc = blo_occupancy<2>(&occ); break;
    case 4: ea = cudaFuncSetAttribute(blo_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_ring_bytes<4>()); rc = blo_occupancy<4>(&occ); break;
    default: ea = cudaFuncSetAttribute(blo_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_ring_bytes<8>()); rc = blo_occupancy<8>(&occ); break;
  }
  if (ea != cudaSuccess) return fail(PTGPU_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(ea));
  if (rc) return rc;
  if (occ < 1) return fail(PTGPU_ERR_CUDA, "step-program kernel does not fit an SM");
  e->blo_blocks_per_sm = occ;
  return PTGPU_OK;
}

// Every index of the program against the engine's and the candidate's own ranges.
// Local optimisation (procedure 0) may write private slots only; the whole-tree
// procedure works on the engine's own buffers, like the calls it replaces.
int blo_validate(ptgpu_engine *e, const BloRequest &r)
{
  const unsigned int count = r.count, npriv = r.n_private_clvs, nlen = r.n_lengths;
  const unsigned int nsteps = r.n_prologue + r.n_bodies * r.n_body;
  const unsigned int clv_total = e->tips + e->clv_buffers;
  const bool in_place = r.procedure == 1;
  auto clv_ok = [&](unsigned int idx, unsigned int i, bool writable) {
    if (idx < clv_total) return !writable || (in_place && idx >= e->tips);
    return idx - clv_total >= i * npriv && idx - clv_total < (i + 1) * npriv;
  };
  auto scal_ok = [&](int idx, unsigned int i, bool writable) {
    if (idx == PTGPU_SCALE_BUFFER_NONE) return true;
    if (idx < 0) return false;
    if ((unsigned int)idx < e->scale_buffers) return !writable || in_place;
    const unsigned int off = (unsigned int)idx - e->scale_buffers;
    return off >= i * npriv && off < (i + 1) * npriv;
  };
  auto mat_ok = [&](unsigned int idx, unsigned int i, bool writable) {
    if (idx < e->prob_matrices) return !writable || in_place;
    if (in_place) return false;
    const unsigned int off = idx - e->prob_matrices;
    return off >= i * nlen && off < (i + 1) * nlen;
  };
  // the matrix an optimised / recomputed edge owns: its private slot, or (in place) the
  // engine matrix with the slot's number
  auto owns = [&](unsigned int mat, unsigned int slot, unsigned int i) {
    if (slot >= nlen) return false;
    return in_place ? mat == slot && mat < e->prob_matrices : mat == e->prob_matrices + i * nlen + slot;
  };
  for (unsigned int k = 0; k < nsteps; ++k)
    for (unsigned int i = 0; i < count; ++i)
    {
      const ptgpu_blo_step_t &st = r.steps[(size_t)k * count + i];
      if (st.type == PTGPU_BLO_NOP) continue;
      if (st.type != r.step_types[k])
        return fail(PTGPU_ERR_ARG, "step %u of candidate %u has type %d, expected %d", k, i, st.type, r.step_types[k]);
      bool ok;
      if (st.type == PTGPU_BLO_OP)
        ok = clv_ok(st.a_clv, i, true) && scal_ok(st.a_scaler, i, true) && clv_ok(st.b_clv, i, false) &&
             clv_ok(st.c_clv, i, false) && scal_ok(st.b_scaler, i, false) && scal_ok(st.c_scaler, i, false) &&
             mat_ok(st.b_matrix, i, false) && mat_ok(st.c_matrix, i, false) && st.a_clv != st.b_clv &&
             st.a_clv != st.c_clv;
      else if (st.type == PTGPU_BLO_OPT)
        ok = clv_ok(st.a_clv, i, false) && clv_ok(st.b_clv, i, false) && owns(st.b_matrix, st.length_slot, i);
      else if (st.type == PTGPU_BLO_MAT)
        ok = owns(st.b_matrix, st.length_slot, i);
      else
        ok = false;
      if (!ok) return fail(PTGPU_ERR_ARG, "step %u of candidate %u: index out of range", k, i);
    }
  for (unsigned int i = 0; i < count; ++i)
    for (const ptgpu_blo_edge_t *ed : {&r.smooth_edges[i], &r.final_edges[i]})
      if (!clv_ok(ed->parent_clv, i, false) || !clv_ok(ed->child_clv, i, false) ||
          !scal_ok(ed->parent_scaler, i, false) || !scal_ok(ed->child_scaler, i, false) ||
          !mat_ok(ed->matrix, i, false))
        return fail(PTGPU_ERR_ARG, "edge of candidate %u: index out of range", i);
  for (size_t i = 0; i < (size_t)count * nlen; ++i)
    if (!(r.lengths0[i] >= 0)) return fail(PTGPU_ERR_ARG, "bad start length");
  for (unsigned int k = 0; k < r.n_prior; ++k)
    if (r.prior_slots[k] >= nlen) return fail(PTGPU_ERR_ARG, "prior slot out of range");
  return PTGPU_OK;
}

int run_blo(ptgpu_engine *e, const BloRequest &r)
{
  int rc;
  // site shards: the kernel sums over the shards itself (peer memory); a caller-supplied hook
  // cannot be called from inside a kernel
  if (e->allreduce)
    return fail(PTGPU_ERR_STATE, "step programs on a site shard need the engines' own exchange "
                                 "(ptgpu_comm_init / ptgpu_shard_connect), not an allreduce hook");
  if (e->comm.nranks > 1 && (size_t)r.count * 2 > kShardSlots)
    return fail(PTGPU_ERR_ARG, "site shards: at most %u candidates per step program", kShardSlots / 2);
  if ((rc = blo_validate(e, r))) return rc;
  CU(cudaSetDevice(e->device));
  if ((rc = blo_prepare_kernel(e))) return rc;

  const unsigned int count = r.count, npriv = r.n_private_clvs, nlen = r.n_lengths;
  const unsigned int nsteps = r.n_prologue + r.n_bodies * r.n_body;
  const unsigned int clv_total = e->tips + e->clv_buffers;
  const bool in_place = r.procedure == 1;

  // ---- scratch: private CLVs / scalers / matrices / sumtables for a chunk of candidates
  const size_t stride = e->dev.clv_stride;
  const size_t clv_bytes = stride * sizeof(double) + e->dev.scaler_stride * sizeof(unsigned int);
  const size_t per_cand = (size_t)npriv * clv_bytes + stride * sizeof(double) +
                          (size_t)(e->nchunks * 3 + e->ntiles) * sizeof(double) +
                          (size_t)nlen * (e->cats * 16 + 1) * sizeof(double);
  size_t chunk = count;
  // PTGPU_BLO_MAX_CHUNK: candidates per launch (tests use it to exercise the chunked path that
  // otherwise needs a batch larger than the free memory)
  if (e->blo_max_chunk && chunk > e->blo_max_chunk) chunk = e->blo_max_chunk;
  const bool fits = chunk <= e->lcap_cand && npriv <= e->lcap_clvs && nlen <= e->lcap_lens;
  if (!fits)
  {
    // grow the scratch: as many candidates per launch as 80 % of the free memory holds
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    const size_t have = e->lcap_cand * ((size_t)e->lcap_clvs * clv_bytes + stride * sizeof(double));
    const size_t can = std::max<size_t>(1, (size_t)((free_b + have) * 0.8) / per_cand);
    // all shards must cut a batch into the same launches; free memory may differ between them
    if (e->comm.nranks > 1 && can < chunk)
      return fail(PTGPU_ERR_STATE, "site shards: a batch of %zu candidates does not fit this shard's memory "
                                   "(set PTGPU_BLO_MAX_CHUNK to the same value on every shard)", chunk);
    chunk = std::min<size_t>(chunk, can);
  }
  if (!fits && (chunk > e->lcap_cand || npriv > e->lcap_clvs || nlen > e->lcap_lens))
  {
    const size_t nc = std::max<size_t>(chunk, e->lcap_cand), ncl = std::max<size_t>(npriv, e->lcap_clvs),
                 nl = std::max<size_t>(nlen, e->lcap_lens);
    PT_FREE(e, e->d_lclv); PT_FREE(e, e->d_lscal); PT_FREE(e, e->d_lmat); PT_FREE(e, e->d_lsum); PT_FREE(e, e->d_llens);
    PT_FREE(e, e->d_lpart); PT_FREE(e, e->d_ltile); PT_FREE(e, e->d_ledges); PT_FREE(e, e->d_lcand); PT_FREE(e, e->d_lnst);
    PT_FREE(e, e->d_lfin);
    e->d_lclv = e->d_lsum = e->d_lmat = e->d_llens = e->d_lpart = e->d_ltile = nullptr;
    e->d_lscal = nullptr; e->d_ledges = nullptr; e->d_lcand = nullptr; e->d_lnst = nullptr; e->d_lfin = nullptr;
    e->lcap_cand = e->lcap_clvs = e->lcap_lens = 0;
    CU(PT_MALLOC(e, &e->d_lclv, std::max<size_t>(1, nc * ncl) * stride * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_lscal, std::max<size_t>(1, nc * ncl) * e->dev.scaler_stride * sizeof(unsigned int)));
    CU(PT_MALLOC(e, &e->d_lmat, nc * nl * e->cats * 16 * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_lsum, nc * stride * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_llens, nc * nl * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_lpart, 2 * nc * e->nchunks * 3 * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_ltile, nc * e->ntiles * sizeof(double)));
    CU(PT_MALLOC(e, &e->d_ledges, nc * 2 * sizeof(BloEdge)));
    CU(PT_MALLOC(e, &e->d_lcand, nc * sizeof(BloCand)));
    CU(PT_MALLOC(e, &e->d_lnst, 2 * nc * sizeof(NewtonState)));
    CU(PT_MALLOC(e, &e->d_lfin, nc * sizeof(int)));
    e->lcap_cand = nc; e->lcap_clvs = ncl; e->lcap_lens = nl;
  }
  if ((rc = ensure(&e->d_lsteps, &e->lcap_steps, (size_t)nsteps * chunk))) return rc;
  if ((rc = ensure(&e->d_ltypes, &e->lcap_types, (size_t)nsteps))) return rc;
  if ((rc = ensure(&e->d_lprior, &e->lcap_prior, (size_t)std::max(1u, r.n_prior)))) return rc;
  // descriptors and results travel through the engine's pinned arena (see ptgpu_engine::h_pin)
  if ((rc = pin_reserve(e, PinCursor::padded(nsteps * sizeof(int)) +
                               PinCursor::padded(std::max(1u, r.n_prior) * sizeof(unsigned int)) +
                               PinCursor::padded((size_t)nsteps * chunk * sizeof(BloStep)) +
                               PinCursor::padded(2 * chunk * sizeof(BloEdge)) +
                               PinCursor::padded(chunk * sizeof(BloCand)) +
                               PinCursor::padded(chunk * nlen * sizeof(double)))))
    return rc;
  CU(cudaStreamSynchronize(e->stream));  // nothing in flight still reads the descriptor buffers
  PinCursor pin{e->h_pin};
  int *htypes = pin.take<int>(nsteps);
  unsigned int *hprior = pin.take<unsigned int>(std::max(1u, r.n_prior));
  BloStep *hsteps = pin.take<BloStep>((size_t)nsteps * chunk);
  BloEdge *hedges = pin.take<BloEdge>(2 * chunk);
  BloCand *hcand = pin.take<BloCand>(chunk);
  double *hlens = pin.take<double>(chunk * nlen);
  memcpy(htypes, r.step_types, nsteps * sizeof(int));
  CU(cudaMemcpyAsync(e->d_ltypes, htypes, nsteps * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  if (r.n_prior)
  {
    memcpy(hprior, r.prior_slots, r.n_prior * sizeof(unsigned int));
    CU(cudaMemcpyAsync(e->d_lprior, hprior, r.n_prior * sizeof(unsigned int), cudaMemcpyHostToDevice, e->stream));
  }

  for (unsigned int base = 0; base < count; base += (unsigned int)chunk)
  {
    const unsigned int nb = (unsigned int)std::min<size_t>(chunk, count - base);
    // re-pack this chunk's columns; private indices are rebased to the chunk
    const unsigned int dclv = base * npriv, dmat = base * nlen;
    auto rb_clv = [&](unsigned int idx) { return idx >= clv_total ? idx - dclv : idx; };
    auto rb_scal = [&](int idx) { return (idx >= 0 && (unsigned int)idx >= e->scale_buffers) ? (int)(idx - dclv) : idx; };
    auto rb_mat = [&](unsigned int idx) { return idx >= e->prob_matrices ? idx - dmat : idx; };
    for (unsigned int k = 0; k < nsteps; ++k)
      for (unsigned int i = 0; i < nb; ++i)
      {
        const ptgpu_blo_step_t &st = r.steps[(size_t)k * count + base + i];
        BloStep d{};
        d.type = st.type;
        d.a_clv = rb_clv(st.a_clv); d.a_scal = rb_scal(st.a_scaler);
        d.b_clv = rb_clv(st.b_clv); d.b_mat = rb_mat(st.b_matrix); d.b_scal = rb_scal(st.b_scaler);
        d.c_clv = rb_clv(st.c_clv); d.c_mat = rb_mat(st.c_matrix); d.c_scal = rb_scal(st.c_scaler);
        d.slot = st.length_slot;
        hsteps[(size_t)k * nb + i] = d;
      }
    for (unsigned int i = 0; i < nb; ++i)
      for (int w = 0; w < 2; ++w)
      {
        const ptgpu_blo_edge_t &ed = w ? r.final_edges[base + i] : r.smooth_edges[base + i];
        BloEdge d{};
        d.p_clv = rb_clv(ed.parent_clv); d.p_scal = rb_scal(ed.parent_scaler);
        d.c_clv = rb_clv(ed.child_clv); d.c_scal = rb_scal(ed.child_scaler);
        d.mat = rb_mat(ed.matrix);
        hedges[(size_t)w * nb + i] = d;
      }
    for (unsigned int i = 0; i < nb; ++i)
    {
      hcand[i] = BloCand{};
      hcand[i].active = 1;
    }
    const size_t nlens = (size_t)nb * nlen;
    memcpy(hlens, r.lengths0 + (size_t)base * nlen, nlens * sizeof(double));
    CU(cudaMemcpyAsync(e->d_lsteps, hsteps, (size_t)nsteps * nb * sizeof(BloStep), cudaMemcpyHostToDevice, e->stream));
    CU(cudaMemcpyAsync(e->d_ledges, hedges, (size_t)2 * nb * sizeof(BloEdge), cudaMemcpyHostToDevice, e->stream));
    CU(cudaMemcpyAsync(e->d_lcand, hcand, nb * sizeof(BloCand), cudaMemcpyHostToDevice, e->stream));
    CU(cudaMemcpyAsync(e->d_llens, hlens, nlens * sizeof(double), cudaMemcpyHostToDevice, e->stream));

    BloArgs a{};
    a.v.s = e->dev;
    a.v.clv_total = clv_total; a.v.scal_total = e->scale_buffers; a.v.mat_total = e->prob_matrices;
    a.v.pclv = e->d_lclv; a.v.pscal = e->d_lscal;
    a.v.pmat = in_place ? nullptr : e->d_lmat;   // in place: the engine's own matrices, untouched until a step writes them
    a.m = e->model;
    a.steps = e->d_lsteps; a.step_type = e->d_ltypes;
    a.smooth_edge = e->d_ledges; a.final_edge = e->d_ledges + nb;
    a.cand = e->d_lcand; a.lens = e->d_llens; a.nst = e->d_lnst; a.fin = e->d_lfin;
    a.sumtables = e->d_lsum; a.entry_stride = stride;
    a.partials = e->d_lpart; a.tile_partials = e->d_ltile;
    a.B = nb; a.NL = nlen; a.n_pro = r.n_prologue; a.n_body = r.n_body; a.n_bodies = r.n_bodies;
    a.sites = e->sites; a.ntiles = e->ntiles; a.chunk_tiles = e->chunk_tiles; a.nchunks = e->nchunks;
    a.smoothings = r.smoothings; a.max_newton = r.max_newton;
    a.xmin = r.xmin; a.xmax = r.xmax; a.xtol = r.xtol; a.tolerance = r.tolerance;
    a.vec_end = e->vec_end;
    a.procedure = r.procedure; a.lambda = r.lambda;
    a.prior_slots = e->d_lprior; a.n_prior = r.n_prior;
    a.comm = e->comm;
    const unsigned int work = std::max(1u, nb * e->ntiles);
    const unsigned int grid = std::min(work, coop_grid_share(e, (unsigned int)(e->blo_blocks_per_sm * e->sm_count)));
    // a program whose phases are a couple of tiles per CTA is all barriers: one cluster
    if (work <= e->blo_cluster_work && !PT_SWEEP_RING)
    {
      DISPATCH_CATS(e, rc = launch_blo_cluster<C>(e, a));
    }
    else
    {
      DISPATCH_CATS(e, rc = launch_blo<C>(e, a, grid));
    }
    if (rc) return rc;
    CU(cudaMemcpyAsync(hcand, e->d_lcand, nb * sizeof(BloCand), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaMemcpyAsync(hlens, e->d_llens, nlens * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    if ((rc = shard_check(e))) return rc;
    for (unsigned int i = 0; i < nb; ++i)
    {
      r.lnl_out[base + i] = hcand[i].final_lnl;
      if (r.status_out) r.status_out[base + i] = hcand[i].failed;
      if (r.smoothings_out) r.smoothings_out[base + i] = hcand[i].smoothings;
      e->counters[2] += (unsigned long long)hcand[i].newton_evals;
    }
    if (r.lengths_out)
      memcpy(r.lengths_out + (size_t)base * nlen, hlens, nlens * sizeof(double));
    if (!in_place) e->counters[3] += nb;
  }
  return PTGPU_OK;
}

} // namespace

extern "C" int ptgpu_optimize_local_batch(ptgpu_engine *e, unsigned int count,
                                          unsigned int n_private_clvs, unsigned int n_lengths,
                                          unsigned int n_prologue, unsigned int n_body,
                                          const int *step_types, const ptgpu_blo_step_t *steps,
                                          const ptgpu_blo_edge_t *smooth_edges,
                                          const ptgpu_blo_edge_t *final_edges, const double *lengths0,
                                          double bl_min, double bl_max, double tolerance,
                                          unsigned int smoothings, double *lengths_out,
                                          double *lnl_out, int *status_out, int *smoothings_out)
{
  if (!e || !step_types || !steps || !smooth_edges || !final_edges || !lengths0 || !lnl_out)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_local_batch: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  if (!count) return PTGPU_OK;
  if (!n_private_clvs || !n_lengths || !smoothings)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_local_batch: empty program");
  BloRequest r{};
  r.count = count; r.n_private_clvs = n_private_clvs; r.n_lengths = n_lengths;
  r.n_prologue = n_prologue; r.n_body = n_body; r.n_bodies = 2;
  r.step_types = step_types; r.steps = steps;
  r.smooth_edges = smooth_edges; r.final_edges = final_edges; r.lengths0 = lengths0;
  r.procedure = 0;
  // pll_optimize.c:1011-1021
  r.xmin = bl_min > 0 ? bl_min : 1.0e-4;
  r.xmax = bl_max > 0 ? bl_max : 100.0;
  r.xtol = bl_min > 0 ? bl_min / 10.0 : 1.0e-4;
  r.tolerance = tolerance; r.lambda = 0.0;
  r.smoothings = smoothings; r.max_newton = 30;
  r.lengths_out = lengths_out; r.lnl_out = lnl_out; r.status_out = status_out;
  r.smoothings_out = smoothings_out;
  return run_blo(e, r);
}

extern "C" int ptgpu_optimize_all_branches(ptgpu_engine *e, unsigned int n_prologue,
                                           unsigned int n_body, unsigned int n_bodies,
                                           const ptgpu_blo_step_t *steps,
                                           const ptgpu_blo_edge_t *lnl_edge, const double *lengths0,
                                           const unsigned int *prior_edges, unsigned int n_prior,
                                           unsigned int max_iter, double epsilon, double map_lambda,
                                           unsigned int max_passes, double *lengths_out,
                                           double *lnl_out, int *status_out, int *passes_out)
{
  if (!e || !steps || !lnl_edge || !lengths0 || !lengths_out || !lnl_out)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_all_branches: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  if (!n_body || !n_bodies || !max_passes || !max_iter)
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_all_branches: empty program");
  if (n_prior && (!prior_edges || !(map_lambda > 0)))
    return fail(PTGPU_ERR_ARG, "ptgpu_optimize_all_branches: prior edges without a prior rate");
  const unsigned int nsteps = n_prologue + n_bodies * n_body;
  std::vector<int> types(nsteps);
  for (unsigned int k = 0; k < nsteps; ++k) types[k] = steps[k].type;
  BloRequest r{};
  r.count = 1; r.n_private_clvs = 0; r.n_lengths = e->prob_matrices;
  r.n_prologue = n_prologue; r.n_body = n_body; r.n_bodies = n_bodies;
  r.step_types = types.data(); r.steps = steps;
  r.smooth_edges = lnl_edge; r.final_edges = lnl_edge; r.lengths0 = lengths0;
  r.procedure = 1;
  r.xmin = 0.0; r.xmax = 0.0; r.xtol = epsilon; r.tolerance = epsilon; r.lambda = map_lambda;
  r.smoothings = max_passes; r.max_newton = max_iter;
  r.prior_slots = prior_edges; r.n_prior = n_prior;
  r.lengths_out = lengths_out; r.lnl_out = lnl_out; r.status_out = status_out;
  r.smoothings_out = passes_out;
  int rc = run_blo(e, r);
  if (rc) return rc;
  if (getenv("PTGPU_DEBUG"))
  {
    unsigned int cnt[4] = {0, 0, 0, 0};
    for (unsigned int k = n_prologue; k < n_prologue + n_body; ++k) cnt[steps[k].type & 3] += 1;
    fprintf(stderr, "[ptgpu] all-branches program: prologue %u, body %u x %u (first body: %u OP, %u OPT, %u MAT, %u NOP), "
                    "%d passes, lnL %.6f\n", n_prologue, n_body, n_bodies, cnt[1], cnt[2], cnt[3], cnt[0],
            passes_out ? *passes_out : -1, *lnl_out);
  }
  if (status_out && *status_out)
    return fail(PTGPU_ERR_NOT_FINITE, *status_out == 1 ? "d1 is not finite" : "d2 is not finite");
  return PTGPU_OK;
}

// ---- an entry point that allocates nothing --------------------------------------------------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  ptgpu_compute_edge_loglikelihood's kernel needs scratch (one partial per tile of
// 256 sites, the arrival counter of its fixed-order finish, the result): here that scratch is a
// file-scope device array of a fixed, modest size that every GPU loading this library holds until
// exit.  The array's NAME is only a host-side stand-in in host code; the address a kernel can use
// comes from cudaGetSymbolAddress, taken once per device with that device current.
namespace {

constexpr unsigned int kNoAllocMaxTiles = 8192;                        // 8192 x 256 = 2,097,152 sites
constexpr unsigned int kNoAllocMaxSites = kNoAllocMaxTiles * kTileSites;
// [0, kNoAllocMaxTiles): tile partials; [kNoAllocMaxTiles]: the result; [kNoAllocMaxTiles + 1]: the counter word
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxTiles + 2];

std::mutex g_noalloc_mutex;
double *g_noalloc_ptr[64] = {};   // per device: device address of g_noalloc_scratch

int noalloc_setup(int device, double **out)
{
  if (device < 0 || device >= 64) return fail(PTGPU_ERR_ARG, "no-alloc entry point: device %d out of range", device);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device])
  {
    void *p = nullptr;
    CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // the current device's copy of the array
    g_noalloc_ptr[device] = static_cast<double *>(p);
  }
  *out = g_noalloc_ptr[device];
  return PTGPU_OK;
}

} // namespace

extern "C" unsigned int ptgpu_noalloc_max_sites(void) { return kNoAllocMaxSites; }

extern "C" int ptgpu_compute_edge_loglikelihood_noalloc(ptgpu_engine *e, unsigned int parent_clv_index,
                                                        int parent_scaler_index, unsigned int child_clv_index,
                                                        int child_scaler_index, unsigned int matrix_index,
                                                        double *lnl_out)
{
  if (!e || !lnl_out) return fail(PTGPU_ERR_ARG, "ptgpu_compute_edge_loglikelihood_noalloc: null argument");
  if (!e->model_set) return fail(PTGPU_ERR_STATE, "model not set");
  if (e->sites > kNoAllocMaxSites || e->ntiles > kNoAllocMaxTiles)
    return fail(PTGPU_ERR_ARG, "ptgpu_compute_edge_loglikelihood_noalloc: %u sites exceed the fixed scratch "
                               "(at most %u sites); use ptgpu_compute_edge_loglikelihood", e->sites, kNoAllocMaxSites);
  if (is_sharded(e))
    return fail(PTGPU_ERR_STATE, "ptgpu_compute_edge_loglikelihood_noalloc: not available on a site shard");
  int rc;
  if ((rc = check_clv(e, parent_clv_index, "edge_lnl.parent"))) return rc;
  if ((rc = check_clv(e, child_clv_index, "edge_lnl.child"))) return rc;
  if ((rc = check_scaler(e, parent_scaler_index, "edge_lnl.parent"))) return rc;
  if ((rc = check_scaler(e, child_scaler_index, "edge_lnl.child"))) return rc;
  if ((rc = check_matrix(e, matrix_index, "edge_lnl"))) return rc;
  CU(cudaSetDevice(e->device));
  double *scratch = nullptr;
  if ((rc = noalloc_setup(e->device, &scratch))) return rc;
  double *partials = scratch, *result = scratch + kNoAllocMaxTiles;
  unsigned int *counter = reinterpret_cast<unsigned int *>(scratch + kNoAllocMaxTiles + 1);
  // the array keeps the last call's contents: the arrival counter must start at zero (the kernel
  // re-arms it, but a call that was cut short would not have); partials and result are overwritten
  CU(cudaMemsetAsync(counter, 0, sizeof(double), e->stream));
  e->counters[0]++;
  e->kernel_launches[PTGPU_KERNEL_EDGE_LNL]++;
  DISPATCH_CATS(e, (edge_lnl_kernel<C><<<e->ntiles, kSitesPerPass * C, 0, e->stream>>>(
                       e->dev, e->model, parent_clv_index, parent_scaler_index, child_clv_index,
                       child_scaler_index, matrix_index, nullptr, partials, counter, result)));
  CU(cudaGetLastError());
  // the engine's pinned staging page (allocated when the engine was created) receives the result
  CU(cudaMemcpyAsync(e->stg.host + kStgResult, result, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  *lnl_out = *(volatile double *)(e->stg.host + kStgResult);
  return PTGPU_OK;
}
