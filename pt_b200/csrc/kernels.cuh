// pt_b200/csrc/kernels.cuh -- sm_100a device code of the likelihood engine.
//
// Work decomposition shared by every kernel here: one thread owns one
// (site, rate-category) pair, i.e. one 4-double (32-byte) CLV entry.  A CTA
// covers 64 consecutive sites x C categories per pass and PASSES passes per
// tile, so a warp reads 32 consecutive 32-byte entries = 1 KB contiguous with
// one LDG.E.256 per thread (fully coalesced), and the C lanes of one site sit
// next to each other in a warp so per-site reductions over categories are
// width-C shuffles.  All arithmetic that must match libpll's AVX path bit for
// bit uses __dmul_rn/__dadd_rn (never contracted to FMA) in the reference's
// summation order (note N1).
//
// Reference citations: LIBPLL = lib/libptpll/lib/pll-modules/libs/libpll/src,
// PTPLL = lib/libptpll/src (under /root/reference).
#pragma once

#include <cooperative_groups.h>
#include <cstdint>
#include <cuda_runtime.h>

namespace ptgpu {

namespace cg = cooperative_groups;

constexpr int kSitesPerPass = 64;  // sites covered by a CTA in one pass
constexpr int kPasses = 4;         // passes per tile
constexpr int kTileSites = kSitesPerPass * kPasses;  // 256 sites per tile
constexpr int kMaxCats = 8;
// The two CLV-producing kernels keep P matrices in registers (32 doubles per
// thread) and amortise them over kHeavyPasses sites, so they run smaller CTAs.
constexpr int kHeavySites = 32;   // sites covered by a CTA in one pass
constexpr int kHeavyPasses = 4;   // passes per tile
constexpr int kHeavyTile = kHeavySites * kHeavyPasses; // 128 sites per tile

// Device view of one engine (passed by value; lives in the constant bank).
struct DevState
{
  const uint8_t *tips;   // [n_tips][sites]
  double *clv;           // [clv_buffers][sites][C][4]
  unsigned int *scalers; // [scale_buffers][sites]
  double *pmat;          // [prob_matrices][C][16]
  unsigned int n_tips;
  unsigned int sites;
  size_t clv_stride;     // doubles per CLV buffer
  size_t tip_stride;     // bytes per tip row (sites rounded up to 16: TMA bulk alignment)
  size_t scaler_stride;  // words per scaler row (sites rounded up to 4)
};

// Model constants (kernel parameter => constant bank, broadcast reads).
struct ModelConst
{
  double evecs[16];     // eigenvecs[j*4+k]                  (LIBPLL/models.c:290-318)
  double inv_evecs[16];
  double tt_inv[16];    // tt_inv[j*4+k] = inv_evecs[k*4+j]*freqs[k] (LIBPLL/core_derivatives_avx.c:86-95)
  double evals[4];
  double freqs[4];
  double rates[kMaxCats];
  double rate_weight;   // 1/C                               (LIBPLL/pll.c:742-748)
  double log_scale_thr; // host log(2^-256)                  (LIBPLL/core_likelihood_avx.c:1250)
  // pattern weights [sites] (LIBPLL/pll.c: pll_set_pattern_weights), or null = all 1 as under
  // pt, which never compresses; read only where a per-site value enters a sum over sites
  const unsigned int *weights;
};

struct DevOp // == ptgpu_operation_t
{
  unsigned int parent_clv;
  int parent_scaler;
  unsigned int c1_clv;
  unsigned int c1_mat;
  int c1_scaler;
  unsigned int c2_clv;
  unsigned int c2_mat;
  int c2_scaler;
};

struct DevCand // == ptgpu_nni_candidate_t
{
  DevOp p;
  DevOp q;
  double length0;
};

// A group of NNI candidates that read the same four neighbour CLVs: candidate 0
// combines children (0,1) on the p side and (2,3) on the q side; an optional
// candidate 1 (the other NNI across the same edge) combines (3,1) and (2,0).
struct DevChild
{
  unsigned int clv;
  int scaler;
  unsigned int mat;
  unsigned int pad;
};
struct DevGroup
{
  DevChild ch[4];
  unsigned int ncand;   // 1 or 2
  unsigned int out[2];  // candidate slots (sumtable / scaler / state index)
  unsigned int pad;
};

// Newton state of one branch (PTPLL/pll_partition.cpp:419-481)
struct NewtonState
{
  double len;
  double d1;
  double d2;
  double lnl;
  int iters;
  int done;             // 0 running, 1 finished
  int maybe_decreasing;
  int status;           // mode 0: 0 ok, 1 d1 not finite, 2 d2 not finite
                        // mode 1: 0 ok, 1 derivative not finite, 2 iteration limit
  double xl, xh;        // mode 1: current bracket
  int lnl_valid;        // lnl holds the log-likelihood at `len` (the last sweep was at the final length)
  int try_lnl;          // the next sweep should also sum site log-likelihoods (it is probably the last)
  int lnl_sweeps;       // statistics: sweeps that carried the lnL sums / extra final sweeps
  int final_sweeps;
};

// ---------------------------------------------------------------- site shards
// The only exchange between the GPUs that hold the site shards of one alignment is the sum
// of per-tree scalars: 1 double per edge log-likelihood, 2 (d1, d2) per Newton step -- the
// shape of pll-modules' parallel_reduce_cb hooks (PLLMOD/optimize/pll_optimize.c:1211-1212,
// :1249-1255).  It is done INSIDE the kernels over NVLink peer memory: every rank owns a
// mailbox that all ranks have mapped; a value is published by storing it into every rank's
// mailbox and collected by polling the local one.  Each double travels as two 8-byte words
// {32 data bits, 32-bit sequence flag} (8-byte stores are single-copy atomic, so a word is
// valid exactly when it carries the flag of the exchange the reader waits for): no fence, no
// ordering between words, no host, no NCCL call.  Sums are formed in RANK ORDER by every
// reader, so all ranks get the same bits and results are reproducible per shard count.
// Slots alternate between two halves by exchange parity; a rank publishes exchange q+1 only
// after all its readers of exchange q have passed a grid barrier (or a kernel boundary), and
// a peer can publish q+2 only after it has collected this rank's q+1 -- nobody overwrites a
// word that is still being waited for.
constexpr int kMaxShards = 8;
constexpr unsigned int kShardSlots = 4096;   // doubles per (parity, source rank)

struct ShardComm
{
  int rank, nranks;                      // nranks <= 1: not sharded
  unsigned long long *box[kMaxShards];   // box[r]: rank r's mailbox as mapped here, [2][nranks][kShardSlots][2] words
  unsigned long long *seq;               // this device: exchanges completed so far (same on all ranks)
  int *error;                            // this device: sticky; a peer did not answer within timeout_ns
  unsigned long long timeout_ns;
};

constexpr size_t shard_mailbox_bytes() { return (size_t)2 * kMaxShards * kShardSlots * 2 * sizeof(unsigned long long); }

__device__ __forceinline__ unsigned long long shard_flag(unsigned long long q)
{
  return q % 0xffffffffull + 1ull;   // never 0: a zeroed mailbox matches no exchange
}

__device__ __forceinline__ unsigned long long global_timer_ns()
{
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// publish value `slot` of exchange q to every rank (own mailbox included)
__device__ __forceinline__ void shard_put(const ShardComm &c, unsigned long long q, unsigned int slot, double v)
{
  const unsigned long long bits = (unsigned long long)__double_as_longlong(v), f = shard_flag(q);
  const unsigned long long w0 = (bits << 32) | f, w1 = (bits & 0xffffffff00000000ull) | f;
  const size_t off = (((q & 1ull) * (size_t)c.nranks + (size_t)c.rank) * kShardSlots + slot) * 2;
  for (int d = 0; d < c.nranks; ++d)
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(c.box[d] + off), "l"(w0), "l"(w1) : "memory");
}

// value `slot` of exchange q as published by rank `src` (spins until it has arrived; NaN and
// the sticky error flag if it does not within the timeout, so that a dead peer ends the
// kernel with "derivative not finite" instead of hanging the device)
__device__ __forceinline__ double shard_get(const ShardComm &c, unsigned long long q, unsigned int slot, int src)
{
  const unsigned long long f = shard_flag(q);
  const unsigned long long *p =
      c.box[c.rank] + (((q & 1ull) * (size_t)c.nranks + (size_t)src) * kShardSlots + slot) * 2;
  unsigned long long w0, w1, t0 = 0;
  unsigned int spins = 0;
  for (;;)
  {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
    if ((w0 & 0xffffffffull) == f && (w1 & 0xffffffffull) == f) break;
    if ((++spins & 0xffu) == 0)
    {
      if (*(volatile int *)c.error) return __longlong_as_double(0x7ff8000000000000ll);
      const unsigned long long now = global_timer_ns();
      if (!t0) t0 = now;
      else if (now - t0 > c.timeout_ns)
      {
        *(volatile int *)c.error = 1;
        return __longlong_as_double(0x7ff8000000000000ll);
      }
    }
  }
  return __longlong_as_double((long long)((w0 >> 32) | (w1 & 0xffffffff00000000ull)));
}

// Sums over ranks of values slot0 .. slot0+NV-1 of exchange q, in rank order.  Called by all
// 32 lanes of one (converged) warp: lane r waits for rank r's words, so the polls overlap;
// every lane returns the same sums.
template <int NV>
__device__ __forceinline__ void shard_collect(const ShardComm &c, unsigned long long q, unsigned int slot0,
                                              double (&out)[NV])
{
  const int lane = threadIdx.x & 31;
  double mine[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) mine[v] = lane < c.nranks ? shard_get(c, q, slot0 + v, lane) : 0.0;
  __syncwarp();
#pragma unroll
  for (int v = 0; v < NV; ++v)
  {
    double acc = __shfl_sync(0xffffffffu, mine[v], 0);
    for (int r = 1; r < c.nranks; ++r) acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, mine[v], r));
    out[v] = acc;
  }
}

// ---------------------------------------------------------------- helpers

#define PT_SCALE_FACTOR 1.157920892373161954235709850086879078532699846656405640394575840079131296399360e77 /* 2^256 */
#define PT_SCALE_THRESHOLD (1.0 / PT_SCALE_FACTOR)

__device__ __forceinline__ void ld256(const double *p, double v[4])
{
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
               : "l"(p));
}

// streaming read of data nobody re-reads soon (sumtable sweeps)
__device__ __forceinline__ void ld256_stream(const double *p, double v[4])
{
  asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
               : "l"(p));
}

// read-only path (P-matrix rows: never written while a kernel that uses this runs)
__device__ __forceinline__ void ld256_nc(const double *p, double v[4])
{
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
               : "l"(p));
}

__device__ __forceinline__ void st256(double *p, const double v[4])
{
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]),
               "d"(v[2]), "d"(v[3])
               : "memory");
}

// ---- TMA bulk copies (cp.async.bulk, SASS UBLKCP) completing on an mbarrier ----
__device__ __forceinline__ unsigned int smem_addr(const void *p)
{
  return (unsigned int)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned int bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned int bytes,
                                         unsigned long long *bar)
{
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_addr(dst)),
      "l"(src), "r"(bytes), "r"(smem_addr(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned int parity)
{
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void lds256(const double *p, double v[4])
{
  const double2 a = *reinterpret_cast<const double2 *>(p);
  const double2 b = *reinterpret_cast<const double2 *>(p + 2);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// (a0 b0 + a1 b1) + (a2 b2 + a3 b3), every product and sum rounded separately
// PT_FMA=1 (experiment only, never the shipped build: scripts/fma_variant.sh): the same tree with
// the second product of each pair fused into the addition -- 4 instead of 7 FP64 issue slots, but
// not libpll's bits (SURVEY.md section 7 asked for it to be measured behind the scaler / 1e-9 gates)
__device__ __forceinline__ double dot4(const double *a, const double b[4])
{
#if defined(PT_FMA) && PT_FMA
  return __dadd_rn(__fma_rn(a[0], b[0], __dmul_rn(a[1], b[1])), __fma_rn(a[2], b[2], __dmul_rn(a[3], b[3])));
#else
  return __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])),
                   __dadd_rn(__dmul_rn(a[2], b[2]), __dmul_rn(a[3], b[3])));
#endif
}

// One (site, cat) entry of CLV `clv_index`: tips are expanded from their 4-bit
// mask to the 0.0/1.0 vector libpll stores (LIBPLL/pll.c:920-925).
template <int C>
__device__ __forceinline__ void load_entry(const DevState &s, unsigned int clv_index,
                                           unsigned int site, int cat, double v[4])
{
  if (clv_index < s.n_tips)
  {
    const unsigned int m = s.tips[(size_t)clv_index * s.tip_stride + site];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ((m >> j) & 1u) ? 1.0 : 0.0;
  }
  else
  {
    ld256(s.clv + (size_t)(clv_index - s.n_tips) * s.clv_stride +
              ((size_t)site * C + cat) * 4,
          v);
  }
}

__device__ __forceinline__ unsigned int load_scaler(const DevState &s, int scaler_index,
                                                    unsigned int site)
{
  return scaler_index < 0 ? 0u : s.scalers[(size_t)scaler_index * s.scaler_stride + site];
}

// AND of a predicate over the C lanes that hold one site
template <int C>
__device__ __forceinline__ bool site_all(bool pred)
{
  unsigned int b = pred ? 1u : 0u;
#pragma unroll
  for (int d = 1; d < C; d <<= 1) b &= __shfl_xor_sync(0xffffffffu, b, d, C);
  return b != 0;
}

// x_i = dot4(M[i], child) for the four rows of a staged 4x4 matrix, then the
// elementwise product of the two children's vectors: one libpll CLV entry
// (LIBPLL/core_partials_avx.c:418-476).
__device__ __forceinline__ void combine_entry(const double *lm, const double *rm,
                                              const double l[4], const double r[4],
                                              double p[4])
{
#pragma unroll
  for (int i = 0; i < 4; ++i)
    p[i] = __dmul_rn(dot4(lm + 4 * i, l), dot4(rm + 4 * i, r));
}

// deterministic CTA-wide sum of one double per thread; result valid in thread 0
template <int THREADS>
__device__ __forceinline__ double block_sum(double v, double *smem /* >= THREADS/32 */)
{
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, d));
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) smem[w] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x < 32)
  {
    t = (threadIdx.x < THREADS / 32) ? smem[threadIdx.x] : 0.0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) t = __dadd_rn(t, __shfl_down_sync(0xffffffffu, t, d));
  }
  __syncthreads();
  return t;
}

// N CTA-wide sums at once: the same reduction tree per value as block_sum (so the same bits),
// but one pair of CTA barriers for all N instead of one pair each.  smem >= N * THREADS/32.
// Results valid in thread 0.
template <int THREADS, int N>
__device__ __forceinline__ void block_sum_n(double (&v)[N], double *smem)
{
  constexpr int W = THREADS / 32 < 1 ? 1 : THREADS / 32;
#pragma unroll
  for (int k = 0; k < N; ++k)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v[k] = __dadd_rn(v[k], __shfl_down_sync(0xffffffffu, v[k], d));
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0)
  {
#pragma unroll
    for (int k = 0; k < N; ++k) smem[k * W + w] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < 32)
  {
#pragma unroll
    for (int k = 0; k < N; ++k)
    {
      double t = (threadIdx.x < THREADS / 32) ? smem[k * W + threadIdx.x] : 0.0;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) t = __dadd_rn(t, __shfl_down_sync(0xffffffffu, t, d));
      v[k] = t;
    }
  }
  __syncthreads();
}

// Tip encoding: raw alignment characters -> 4-bit state masks through a
// 256-entry table (pll_map_nt semantics, LIBPLL/maps.c:46-64); the position
// of the first illegal character (if any) is reported through *bad.
struct TipLut { uint8_t m[256]; };
__global__ void tip_encode_kernel(const unsigned char *__restrict__ raw, uint8_t *out,
                                  unsigned int n, TipLut lut, unsigned int *bad)
{
  __shared__ uint8_t t[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) t[i] = lut.m[i];
  __syncthreads();
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const uint8_t mk = t[raw[i]];
    if (!mk) atomicMin(bad, i);
    out[i] = mk;
  }
}

// several tips at once: raw = [count][sites] characters, tip k goes to row tip_of[k]
__global__ void tip_encode_batch_kernel(const unsigned char *__restrict__ raw, uint8_t *tips,
                                        const unsigned int *__restrict__ tip_of, unsigned int count,
                                        unsigned int sites, size_t tip_stride, TipLut lut,
                                        unsigned long long *bad)
{
  __shared__ uint8_t t[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) t[i] = lut.m[i];
  __syncthreads();
  const unsigned long long n = (unsigned long long)count * sites;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const unsigned int k = (unsigned int)(i / sites), site = (unsigned int)(i % sites);
    const uint8_t mk = t[raw[i]];
    if (!mk) atomicMin(bad, i);
    tips[(size_t)tip_of[k] * tip_stride + site] = mk;
  }
}

// ------------------------------------------------------------ kernel 2
// P(t) = I + (V^-1 o expm1(lambda r t)) V per branch and category
// (LIBPLL/core_pmatrix_avx.c:42-310).  One thread per (branch, category).
// lengths/indices live in device memory so that the Newton kernel's result
// can feed this kernel without a host round trip.
// one (branch, category) block of 16 doubles
__device__ __forceinline__ void pmatrix_block(const ModelConst &m, double bl, unsigned int c, double *pm)
{
  if (bl == 0.0)
  {
#pragma unroll
    for (int i = 0; i < 16; ++i) pm[i] = (i % 5 == 0) ? 1.0 : 0.0;
    return;
  }
  double e[4];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    e[k] = expm1(__dmul_rn(__dmul_rn(m.evals[k], m.rates[c]), bl));
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    double w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) w[k] = __dmul_rn(m.inv_evecs[i * 4 + k], e[k]);
    double row[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      const double a = __dadd_rn(__dmul_rn(w[0], m.evecs[0 + j]), __dmul_rn(w[1], m.evecs[4 + j]));
      const double b2 = __dadd_rn(__dmul_rn(w[2], m.evecs[8 + j]), __dmul_rn(w[3], m.evecs[12 + j]));
      row[j] = __dadd_rn(__dadd_rn(a, b2), (i == j) ? 1.0 : 0.0);
    }
    st256(pm + 4 * i, row);
  }
}

template <int C>
__global__ void pmatrix_kernel(DevState s, ModelConst m, const unsigned int *idx,
                               const double *len, unsigned int count)
{
  const unsigned int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count * C) return;
  const unsigned int b = t / C, c = t % C;
  pmatrix_block(m, len[b], c, s.pmat + ((size_t)idx[b] * C + c) * 16);
}

// ------------------------------------------------------------ kernel 1
// Streaming CLV update: every thread walks the whole operation list for the
// (site, category) entries it owns (LIBPLL/partials.c:177-213 runs the same
// list one op at a time over all sites).  Entries of different sites never
// interact, and an entry written by op i is re-read by op j > i only by the
// thread that wrote it, so the kernel has no block-level barrier at all: warps
// free-run through the list.  Because a CTA owns a fixed tile of sites, the
// children an op reads were written by the same SM a few ops earlier and are
// normally still in L2 -- only tips and first touches come from HBM.
// P-matrix rows cost LSU bandwidth like any other load (32 bytes returned to
// every lane), so each thread loads the two 4x4 matrices of its category once
// per op into registers and reuses them for kHeavyPasses sites.
template <int C>
__device__ __forceinline__ void load_matrix(const double *pm, double m[16])
{
#pragma unroll
  for (int i = 0; i < 4; ++i) ld256_nc(pm + 4 * i, m + 4 * i);
}

#ifndef PT_PART_MINB
#define PT_PART_MINB 1
#endif
#ifndef PT_PREP_MINB
#define PT_PREP_MINB 1
#endif
template <int C>
__global__ void __launch_bounds__(kHeavySites *C, PT_PART_MINB)
    partials_kernel(DevState s, const DevOp *__restrict__ ops, unsigned int nops)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = blockIdx.x * kHeavyTile + threadIdx.x / C;

  for (unsigned int o = 0; o < nops; ++o)
  {
    const DevOp op = ops[o];
    double lm[16], rm[16];
    load_matrix<C>(s.pmat + ((size_t)op.c1_mat * C + cat) * 16, lm);
    load_matrix<C>(s.pmat + ((size_t)op.c2_mat * C + cat) * 16, rm);
    double *dst = s.clv + (size_t)(op.parent_clv - s.n_tips) * s.clv_stride;
#pragma unroll 1
    for (int half = 0; half < kHeavyPasses; half += 2)
    {
      double l[2][4], r[2][4], p[2][4];
      unsigned int sc[2];
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int site = min(site0 + (half + j) * kHeavySites, s.sites - 1);
        load_entry<C>(s, op.c1_clv, site, cat, l[j]);
        load_entry<C>(s, op.c2_clv, site, cat, r[j]);
        sc[j] = load_scaler(s, op.c1_scaler, site) + load_scaler(s, op.c2_scaler, site);
      }
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        combine_entry(lm, rm, l[j], r[j], p[j]);
        const unsigned int site = site0 + (half + j) * kHeavySites;
        // per-site scaling (LIBPLL/core_partials_avx.c:491-527): ordered "<"
        bool small = (p[j][0] < PT_SCALE_THRESHOLD) & (p[j][1] < PT_SCALE_THRESHOLD) &
                     (p[j][2] < PT_SCALE_THRESHOLD) & (p[j][3] < PT_SCALE_THRESHOLD);
        small = site_all<C>(small) && op.parent_scaler >= 0;
        if (small)
        {
#pragma unroll
          for (int i = 0; i < 4; ++i) p[j][i] = __dmul_rn(p[j][i], PT_SCALE_FACTOR);
        }
        if (site < s.sites)
        {
          st256(dst + ((size_t)site * C + cat) * 4, p[j]);
          if (cat == 0 && op.parent_scaler >= 0)
            s.scalers[(size_t)op.parent_scaler * s.scaler_stride + site] = sc[j] + (small ? 1u : 0u);
        }
      }
    }
    __syncwarp(); // scaler word written by lane cat 0, re-read by its C-lane group
  }
}

// ---- kernel 1, version 4: P matrices staged per warp in shared memory ----------
// Same barrier-free walk as partials_kernel, but the two 4x4xC matrices of an
// operation are no longer held in 64 registers per thread: every warp keeps a
// private, double-buffered copy in shared memory, filled one operation AHEAD
// with cp.async (LDGSTS) so neither the descriptor->matrix address chain nor
// the matrix fetch sits on the critical path, and consumed row by row
// (LDS.128 broadcast; category blocks are padded to 18 doubles so the four
// category rows of a quarter-warp fall into different banks).  ~80 registers
// instead of 126 => 6 instead of 4 resident CTAs per SM for the same tile.
constexpr int kMatPad = 18;  // doubles per (matrix, category) block in shared memory

__device__ __forceinline__ void cp_async16(void *dst, const void *src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int C>
__device__ __forceinline__ void stage_matrices(const DevState &s, const DevOp &op, double *dst, int lane)
{
  // 2 matrices x C categories x 8 chunks of 16 bytes
#pragma unroll
  for (int c = lane; c < 2 * C * 8; c += 32)
  {
    const int which = c / (C * 8), rest = c % (C * 8), cat = rest / 8, within = rest % 8;
    const double *src = s.pmat + ((size_t)(which ? op.c2_mat : op.c1_mat) * C + cat) * 16 + within * 2;
    cp_async16(dst + (which * C + cat) * kMatPad + within * 2, src);
  }
  cp_async_commit();
}

// PASSES = sites per thread and op: 4 (tiles of 128 sites) on long alignments; 2 (tiles of 64) when the
// alignment has too few 128-site tiles to fill the SMs -- a deep tree on a mid-sized alignment is a
// chain of dependent L2 round trips per warp, and only more warps hide it.  Per-site arithmetic is
// the same, so CLVs and scalers are bit-identical either way.
template <int C, int PASSES = kHeavyPasses>
__global__ void __launch_bounds__(kHeavySites *C, (C <= 4 ? 6 : 1))
    partials_kernel_v4(DevState s, const DevOp *__restrict__ ops, unsigned int nops)
{
  constexpr int WARPS = (kHeavySites * C + 31) / 32;
  __shared__ __align__(16) double mats[WARPS][2][2 * C * kMatPad];
  const int cat = threadIdx.x % C;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int site0 = blockIdx.x * (kHeavySites * PASSES) + threadIdx.x / C;

  DevOp op = ops[0];
  stage_matrices<C>(s, op, mats[warp][0], lane);
  for (unsigned int o = 0; o < nops; ++o)
  {
    const int stage = o & 1;
    DevOp nxt = op;
    if (o + 1 < nops)
    {
      nxt = ops[o + 1];
      stage_matrices<C>(s, nxt, mats[warp][stage ^ 1], lane);
      cp_async_wait<1>();
    }
    else
      cp_async_wait<0>();
    __syncwarp();
    const double *lmat = mats[warp][stage] + cat * kMatPad;
    const double *rmat = mats[warp][stage] + (C + cat) * kMatPad;
    double *dst = s.clv + (size_t)(op.parent_clv - s.n_tips) * s.clv_stride;
#pragma unroll 1
    for (int half = 0; half < PASSES; half += 2)
    {
      double l[2][4], r[2][4], p[2][4];
      unsigned int sc[2];
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int site = min(site0 + (half + j) * kHeavySites, s.sites - 1);
        load_entry<C>(s, op.c1_clv, site, cat, l[j]);
        load_entry<C>(s, op.c2_clv, site, cat, r[j]);
        sc[j] = load_scaler(s, op.c1_scaler, site) + load_scaler(s, op.c2_scaler, site);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
      {
        double lrow[4], rrow[4];
        lds256(lmat + 4 * i, lrow);
        lds256(rmat + 4 * i, rrow);
#pragma unroll
        for (int j = 0; j < 2; ++j) p[j][i] = __dmul_rn(dot4(lrow, l[j]), dot4(rrow, r[j]));
      }
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int site = site0 + (half + j) * kHeavySites;
        bool small = (p[j][0] < PT_SCALE_THRESHOLD) & (p[j][1] < PT_SCALE_THRESHOLD) &
                     (p[j][2] < PT_SCALE_THRESHOLD) & (p[j][3] < PT_SCALE_THRESHOLD);
        small = site_all<C>(small) && op.parent_scaler >= 0;
        if (small)
        {
#pragma unroll
          for (int i = 0; i < 4; ++i) p[j][i] = __dmul_rn(p[j][i], PT_SCALE_FACTOR);
        }
        if (site < s.sites)
        {
          st256(dst + ((size_t)site * C + cat) * 4, p[j]);
          if (cat == 0 && op.parent_scaler >= 0)
            s.scalers[(size_t)op.parent_scaler * s.scaler_stride + site] = sc[j] + (small ? 1u : 0u);
        }
      }
    }
    __syncwarp(); // scaler words / this stage's matrices are free for the op after next
    op = nxt;
  }
}

// Fixed-order finish shared by the reduction kernels: every CTA drops its
// per-tile partial(s) into `partials[tile*NV + v]`; the last CTA to arrive
// sums all tiles in a fixed tree, so the result depends only on `sites`,
// never on launch geometry, SM count or timing.
template <int THREADS, int NV>
__device__ __forceinline__ void finish_tiles(const double (&mine)[NV], double *partials,
                                             unsigned int ntiles, unsigned int *counter,
                                             double *result, double *smem)
{
  __shared__ bool last;
  if (threadIdx.x == 0)
  {
#pragma unroll
    for (int v = 0; v < NV; ++v) partials[(size_t)blockIdx.x * NV + v] = mine[v];
    __threadfence();
    last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
#pragma unroll
  for (int v = 0; v < NV; ++v)
  {
    double acc = 0.0;
    for (unsigned int t = threadIdx.x; t < ntiles; t += THREADS)
      acc = __dadd_rn(acc, __ldcg(partials + (size_t)t * NV + v));
    acc = block_sum<THREADS>(acc, smem);
    if (threadIdx.x == 0) result[v] = acc;
  }
  if (threadIdx.x == 0) *counter = 0u; // re-arm for the next launch on this stream
}

// ------------------------------------------------------------ kernel 3a
// Edge log-likelihood (LIBPLL/core_likelihood_avx.c:1080-1266).
// grid = number of tiles; result[0] = lnL; persite may be null.
template <int C>
__global__ void __launch_bounds__(kSitesPerPass *C)
    edge_lnl_kernel(DevState s, ModelConst m, unsigned int pclv, int pscaler,
                    unsigned int cclv, int cscaler, unsigned int matrix, double *persite,
                    double *partials, unsigned int *counter, double *result)
{
  constexpr int T = kSitesPerPass * C;
  __shared__ double pm[C * 16];
  __shared__ double red[T / 32 < 1 ? 1 : T / 32];
  const int cat = threadIdx.x % C;
  const unsigned int site0 = blockIdx.x * kTileSites + threadIdx.x / C;
  for (int i = threadIdx.x; i < C * 16; i += T) pm[i] = s.pmat[(size_t)matrix * C * 16 + i];
  __syncthreads();

  double p[kPasses][4], c[kPasses][4];
  unsigned int sc[kPasses];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = min(site0 + ps * kSitesPerPass, s.sites - 1);
    load_entry<C>(s, pclv, site, cat, p[ps]);
    load_entry<C>(s, cclv, site, cat, c[ps]);
    sc[ps] = load_scaler(s, pscaler, site) + load_scaler(s, cscaler, site);
  }
  double acc = 0.0;
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = site0 + ps * kSitesPerPass;
    double f[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      f[i] = __dmul_rn(__dmul_rn(m.freqs[i], dot4(pm + cat * 16 + 4 * i, c[ps])), p[ps][i]);
    const double tk = __dadd_rn(__dadd_rn(f[0], f[1]), __dadd_rn(f[2], f[3]));
    // terma += tk * w only for tk > 0, categories in order (:1225-1241)
    double term = (tk > 0.0) ? __dmul_rn(tk, m.rate_weight) : 0.0;
    double terma = term; // lane cat 0 ends up with ((t0 + t1) + t2) + ...
#pragma unroll
    for (int k = 1; k < C; ++k)
      terma = __dadd_rn(terma, __shfl_down_sync(0xffffffffu, term, k, C));
    if (cat == 0 && site < s.sites)
    {
      double lk = log(terma);
      if (sc[ps]) lk = __dadd_rn(lk, __dmul_rn((double)sc[ps], m.log_scale_thr));
      if (m.weights) lk = __dmul_rn(lk, (double)__ldg(m.weights + site));  // :1254
      if (persite) persite[site] = lk;
      acc = __dadd_rn(acc, lk);
    }
  }
  const double mine[1] = {block_sum<T>(acc, red)};
  finish_tiles<T, 1>(mine, partials, gridDim.x, counter, result, red);
}

// ------------------------------------------------------------ kernel 3b
// Sumtable (LIBPLL/core_derivatives_avx.c:25-207): per entry
//   sum_j = dot4(tt_inv[j], clvp) * dot4(evecs[j], clvc)
template <int C>
__device__ __forceinline__ void sumtable_entry(const ModelConst &m, const double p[4],
                                               const double c[4], double out[4])
{
#pragma unroll
  for (int j = 0; j < 4; ++j)
    out[j] = __dmul_rn(dot4(m.tt_inv + 4 * j, p), dot4(m.evecs + 4 * j, c));
}

template <int C>
__global__ void __launch_bounds__(kSitesPerPass *C)
    sumtable_kernel(DevState s, ModelConst m, unsigned int pclv, unsigned int cclv,
                    double *sumtable)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = blockIdx.x * kTileSites + threadIdx.x / C;
  double p[kPasses][4], c[kPasses][4];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = min(site0 + ps * kSitesPerPass, s.sites - 1);
    load_entry<C>(s, pclv, site, cat, p[ps]);
    load_entry<C>(s, cclv, site, cat, c[ps]);
  }
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = site0 + ps * kSitesPerPass;
    double out[4];
    sumtable_entry<C>(m, p[ps], c[ps], out);
    if (site < s.sites) st256(sumtable + ((size_t)site * C + cat) * 4, out);
  }
}

// ------------------------------------------------------------ kernel 4
// Derivatives of -lnL from the sumtable and the Newton-Raphson control loop.

// dg[j][{0,1,2}] = {e, (lambda r) e, (lambda r)^2 e}, e = exp(lambda_j r_c t), for this
// thread's category (LIBPLL/core_derivatives.c:560-575; prop_invar = 0 so
// ki = r_c / 1.0).  Kept in registers: a thread's category never changes.
template <int C>
__device__ __forceinline__ void diag_regs(const ModelConst &m, double len, int cat,
                                          double dg[4][3])
{
  const double ki = m.rates[cat] / (1.0 - 0.0);
#pragma unroll
  for (int j = 0; j < 4; ++j)
  {
    const double lr = __dmul_rn(m.evals[j], ki);
    const double e = exp(__dmul_rn(lr, len));
    dg[j][0] = e;
    dg[j][1] = __dmul_rn(lr, e);
    dg[j][2] = __dmul_rn(__dmul_rn(__dmul_rn(lr, m.evals[j]), ki), e);
  }
}

// Sum over the 4 category lanes of a site for 4 passes at once ("transpose and
// add"): on entry lane k (= category k) holds v[ps][q] for passes ps = 0..3; on
// exit it holds r[q] = sum over categories for pass quad_pass(k).  9 double
// shuffles replace the 36 a per-pass gather would need, and afterwards all
// four lanes (not one) carry a site through the division / log.  The category
// sum is formed as (c0 + c1) + (c2 + c3).
__device__ __forceinline__ int quad_pass(int k) { return ((k & 1) << 1) | (k >> 1); }

template <int NQ>
__device__ __forceinline__ void quad_transpose_sum(const double (&v)[4][NQ], int k,
                                                   double (&r)[NQ])
{
  const bool odd = k & 1, hi = k & 2;
#pragma unroll
  for (int q = 0; q < NQ; ++q)
  {
    const double s0 = odd ? v[0][q] : v[2][q], s1 = odd ? v[1][q] : v[3][q];
    const double k0 = odd ? v[2][q] : v[0][q], k1 = odd ? v[3][q] : v[1][q];
    const double a0 = __dadd_rn(k0, __shfl_xor_sync(0xffffffffu, s0, 1));
    const double a1 = __dadd_rn(k1, __shfl_xor_sync(0xffffffffu, s1, 1));
    const double snd = hi ? a0 : a1, kp = hi ? a1 : a0;
    r[q] = __dadd_rn(kp, __shfl_xor_sync(0xffffffffu, snd, 2));
  }
}

// Sum of per-site log-likelihoods as the log of a running product: every site
// contributes its mantissa to `prod` (one DMUL) and its binary exponent to
// `esum` (integer), and the scaler count enters as -256 per rescaling; one log
// per thread and chunk replaces one log per site.  log(prod) + esum ln 2 equals
// sum_i [log(terma_i) + sc_i log(2^-256)] up to rounding (the product carries
// ~n ulp of relative error, i.e. ~n 1e-16 ABSOLUTE error in the log).
struct LnlProd
{
  double prod = 1.0;
  int esum = 0;
  bool zero = false;   // a site with likelihood 0: the sum is -inf, as the reference's log(0)
  __device__ __forceinline__ void add(double terma, unsigned int sc)
  {
    if (!(terma > 0.0)) { zero = zero || terma == 0.0; if (terma != 0.0) prod = terma; return; }  // NaN propagates
    if (terma < 2.2250738585072014e-308) { terma = __dmul_rn(terma, 1.2676506002282294e30); esum -= 100; }  // denormal: x 2^100
    const int hi = __double2hiint(terma), lo = __double2loint(terma);
    esum += ((hi >> 20) & 0x7ff) - 1022 - 256 * (int)sc;
    prod = __dmul_rn(prod, __hiloint2double((hi & 0x800fffff) | 0x3fe00000, lo));  // mantissa in [0.5, 1)
    if (prod < 3.0549363634996047e-151) { prod = __dmul_rn(prod, 3.2733906078961419e150); esum -= 500; }  // x 2^500
  }
  // pattern of weight w: its likelihood enters the product w times, i.e. mantissa^w (by
  // squaring, renormalised to [0.5, 1) after every product so nothing underflows) and
  // w x exponent; no extra state is carried by the unweighted path
  __device__ __forceinline__ void add_weighted(double terma, unsigned int sc, unsigned int w)
  {
    if (!(terma > 0.0)) { zero = zero || terma == 0.0; if (terma != 0.0) prod = terma; return; }
    int ex = 0;
    if (terma < 2.2250738585072014e-308) { terma = __dmul_rn(terma, 1.2676506002282294e30); ex = -100; }
    const int hi = __double2hiint(terma), lo = __double2loint(terma);
    ex += ((hi >> 20) & 0x7ff) - 1022 - 256 * (int)sc;
    esum += ex * (int)w;
    double b = __hiloint2double((hi & 0x800fffff) | 0x3fe00000, lo), r = 1.0;
    int be = 0, re = 0;
    while (true)
    {
      if (w & 1u)
      {
        r = __dmul_rn(r, b); re += be;
        if (r < 0.5) { r = __dmul_rn(r, 2.0); re -= 1; }
      }
      w >>= 1;
      if (!w) break;
      b = __dmul_rn(b, b); be *= 2;
      if (b < 0.5) { b = __dmul_rn(b, 2.0); be -= 1; }
    }
    esum += re;
    prod = __dmul_rn(prod, r);
    if (prod < 3.0549363634996047e-151) { prod = __dmul_rn(prod, 3.2733906078961419e150); esum -= 500; }
  }
  __device__ __forceinline__ double value() const
  {
    if (zero) return -INFINITY;
    return __dadd_rn(log(prod), __dmul_rn((double)esum, 0.69314718055994530942));
  }
};

// per-site epilogue shared by both reduction layouts
template <bool WITH_LNL>
__device__ __forceinline__ void site_terms(const ModelConst &m, const double lk[3], double terma,
                                           unsigned int site, unsigned int full,
                                           unsigned int sc, double out[3], LnlProd &lp)
{
  // LIBPLL/core_derivatives_avx.c:1111-1156: sites below `full` (= sites/4*4) use
  // the reciprocal form of the 4-site vector step, the tail true divisions
  if (m.weights)
  {
    // pattern weights multiply the site's two terms (:1131-1139, :1153-1154).  All-ones, pt's
    // own case, is m.weights == null and takes the code below untouched.
    const unsigned int w = __ldg(m.weights + site);
    double a, b;
    if (site < full)
    {
      const double recip = __ddiv_rn(1.0, lk[0]);
      a = __dmul_rn(lk[1], recip);
      b = __dadd_rn(__dmul_rn(a, a), -__dmul_rn(lk[2], recip));
      a = -a;
    }
    else
    {
      a = -__ddiv_rn(lk[1], lk[0]);
      b = __dadd_rn(__dmul_rn(a, a), -__ddiv_rn(lk[2], lk[0]));
    }
    out[0] = __dadd_rn(out[0], __dmul_rn(a, (double)w));
    out[1] = __dadd_rn(out[1], __dmul_rn(b, (double)w));
    if (WITH_LNL) lp.add_weighted(terma, sc, w);
    return;
  }
  if (site < full)
  {
    const double recip = __ddiv_rn(1.0, lk[0]);
    const double a = __dmul_rn(lk[1], recip);
    const double b = __dadd_rn(__dmul_rn(a, a), -__dmul_rn(lk[2], recip));
    out[0] = __dadd_rn(out[0], -a);
    out[1] = __dadd_rn(out[1], b);
  }
  else
  {
    const double a = -__ddiv_rn(lk[1], lk[0]);
    const double b = __dadd_rn(__dmul_rn(a, a), -__ddiv_rn(lk[2], lk[0]));
    out[0] = __dadd_rn(out[0], a);
    out[1] = __dadd_rn(out[1], b);
  }
  if (WITH_LNL) lp.add(terma, sc);
}

// One tile of the site sweep.  out[0] += -L'/L terms, out[1] += second
// derivative terms, out[2] += log-likelihood terms (only if WITH_LNL).
// ck[q] of one sumtable entry: ((0 + s0 d0) + s1 d1) + s2 d2) + s3 d3 for q = 0, 1, 2
// (LIBPLL/core_derivatives_avx.c:1019-1034), and the weighted likelihood term if asked for
template <int C, bool WITH_LNL>
__device__ __forceinline__ void sweep_entry_terms(const ModelConst &m, const double (&sv)[4],
                                                  const double (&dg)[4][3], double (&ck)[WITH_LNL ? 4 : 3])
{
#pragma unroll
  for (int q = 0; q < 3; ++q)
  {
    double a = __dadd_rn(0.0, __dmul_rn(sv[0], dg[0][q]));
    a = __dadd_rn(a, __dmul_rn(sv[1], dg[1][q]));
    a = __dadd_rn(a, __dmul_rn(sv[2], dg[2][q]));
    ck[q] = __dadd_rn(a, __dmul_rn(sv[3], dg[3][q]));
  }
  if (WITH_LNL) ck[WITH_LNL ? 3 : 2] = (ck[0] > 0.0) ? __dmul_rn(ck[0], m.rate_weight) : 0.0;
}

// category sums of a tile's entry terms and the per-site contributions to d1, d2 (and lnL)
template <int C, bool WITH_LNL>
__device__ __forceinline__ void sweep_tile_finish(const ModelConst &m, const double (&ck)[kPasses][WITH_LNL ? 4 : 3],
                                                  const unsigned int (&scv)[kPasses], unsigned int sites,
                                                  unsigned int tile, double out[3], LnlProd &lp,
                                                  unsigned long long vec_end)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  // local index of the first site that the reference's scalar tail loop handles
  // (a site shard passes where the GLOBAL sites/4*4 falls inside it)
  const unsigned int full =
      vec_end == ~0ull ? sites / 4 * 4 : (unsigned int)min((unsigned long long)sites, vec_end);
  constexpr int NQ = WITH_LNL ? 4 : 3;
  if (C == 4)
  {
    double r[NQ];
    quad_transpose_sum<NQ>(ck, cat, r);
    const unsigned int site = site0 + quad_pass(cat) * kSitesPerPass;
    if (site < sites) site_terms<WITH_LNL>(m, r, r[NQ - 1], site, full, scv[0], out, lp);
  }
  else
  {
#pragma unroll
    for (int ps = 0; ps < kPasses; ++ps)
    {
      const unsigned int site = site0 + ps * kSitesPerPass;
      double lk[NQ];
#pragma unroll
      for (int q = 0; q < NQ; ++q) lk[q] = ck[ps][q];
#pragma unroll
      for (int k = 1; k < C; ++k)
#pragma unroll
        for (int q = 0; q < NQ; ++q)
          lk[q] = __dadd_rn(lk[q], __shfl_down_sync(0xffffffffu, ck[ps][q], k, C));
      if (cat == 0 && site < sites) site_terms<WITH_LNL>(m, lk, lk[NQ - 1], site, full, scv[ps], out, lp);
    }
  }
}

template <int C, bool WITH_LNL>
__device__ __forceinline__ void sweep_tile_compute(const ModelConst &m, const double (&sv)[kPasses][4],
                                                   const unsigned int (&scv)[kPasses], unsigned int sites,
                                                   unsigned int tile, const double (&dg)[4][3],
                                                   double out[3], LnlProd &lp,
                                                   unsigned long long vec_end)
{
  constexpr int NQ = WITH_LNL ? 4 : 3;
  double ck[kPasses][NQ];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps) sweep_entry_terms<C, WITH_LNL>(m, sv[ps], dg, ck[ps]);
  sweep_tile_finish<C, WITH_LNL>(m, ck, scv, sites, tile, out, lp, vec_end);
}

// The summed scaler words a thread will need in the epilogue of this tile, fetched
// together with the sumtable entries (C == 4: the lane ends up with exactly one
// site, quad_pass(cat); otherwise the category-0 lane handles every pass).
template <int C, bool WITH_LNL>
__device__ __forceinline__ void load_scaler_words(const unsigned int *sclr, unsigned int sites,
                                                  unsigned int tile, unsigned int (&scv)[kPasses])
{
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps) scv[ps] = 0u;
  if (!WITH_LNL || !sclr) return;
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  if (C == 4)
    scv[0] = __ldg(sclr + min(site0 + quad_pass(cat) * kSitesPerPass, sites - 1));
  else if (cat == 0)
  {
#pragma unroll
    for (int ps = 0; ps < kPasses; ++ps) scv[ps] = __ldg(sclr + min(site0 + ps * kSitesPerPass, sites - 1));
  }
}

// direct form: loads straight into registers
template <int C, bool WITH_LNL>
__device__ __forceinline__ void sweep_tile(const ModelConst &m, const double *sumtable,
                                           const unsigned int *sclr, unsigned int sites,
                                           unsigned int tile, const double (&dg)[4][3],
                                           double out[3], LnlProd &lp,
                                           unsigned long long vec_end = ~0ull)
{
  const int cat = threadIdx.x % C;
  const unsigned int site0 = tile * kTileSites + threadIdx.x / C;
  double sv[kPasses][4];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int site = min(site0 + ps * kSitesPerPass, sites - 1);
    ld256_stream(sumtable + ((size_t)site * C + cat) * 4, sv[ps]);
  }
  unsigned int scv[kPasses];
  load_scaler_words<C, WITH_LNL>(sclr, sites, tile, scv);
  sweep_tile_compute<C, WITH_LNL>(m, sv, scv, sites, tile, dg, out, lp, vec_end);
}

// ---- compact sumtables (batched NNI scoring) ------------------------------------------
// A reversible model has one eigenvalue 0 (numerically ~1e-17): column z of the sumtable then
// multiplies exp(0) = 1 in L and (numerically) nothing in L' and L'', whatever the branch
// length.  The batch path stores, per candidate, only the three other columns per (site,
// category) plus ONE double per site, Z = sum over categories of column z -- 13 instead of 16
// doubles per site and Newton evaluation (LIBPLL/core_derivatives.c:560-575 is where the
// reference multiplies by these constants every time).  Layout, per tile of kTileSites sites:
//   [3 planes][kTileSites][C] doubles, then [kTileSites] doubles of Z
// so a warp reads 32 consecutive doubles per plane.  L = (sum over categories of the three-term
// sums) + Z differs from the 16-double form only by the association of one addition per site.
template <int C>
__host__ __device__ constexpr size_t compact_tile_doubles() { return (size_t)kTileSites * (3 * C + 1); }

__device__ __forceinline__ double ld64_stream(const double *p)
{
  double v;
  asm volatile("ld.global.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

template <int C, bool WITH_LNL>
__device__ __forceinline__ void sweep_tile_compact(const ModelConst &m, const double *sumtable,
                                                   const unsigned int *sclr, unsigned int sites,
                                                   unsigned int tile, const double (&dg)[4][3], int zero_col,
                                                   double out[3], LnlProd &lp, unsigned long long vec_end)
{
  constexpr int NQ = 3;
  const int cat = threadIdx.x % C;
  const unsigned int lsite0 = threadIdx.x / C;
  const double *base = sumtable + (size_t)tile * compact_tile_doubles<C>();
  const unsigned int last = sites - 1 - tile * kTileSites;   // local index of the last valid site (if in this tile)
  double sv[kPasses][3];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
    const unsigned int ls = min(lsite0 + ps * kSitesPerPass, last);
#pragma unroll
    for (int p = 0; p < 3; ++p) sv[ps][p] = ld64_stream(base + (size_t)p * kTileSites * C + (size_t)ls * C + cat);
  }
  unsigned int scv[kPasses];
  load_scaler_words<C, WITH_LNL>(sclr, sites, tile, scv);
  // the three live columns' constants, in column order
  double dc[3][3];
#pragma unroll
  for (int p = 0; p < 3; ++p)
  {
    const int j = p + (p >= zero_col ? 1 : 0);
#pragma unroll
    for (int q = 0; q < 3; ++q) dc[p][q] = j == 1 ? dg[1][q] : (j == 2 ? dg[2][q] : (j == 3 ? dg[3][q] : dg[0][q]));
  }
  double ck[kPasses][NQ];
#pragma unroll
  for (int ps = 0; ps < kPasses; ++ps)
  {
#pragma unroll
    for (int q = 0; q < 3; ++q)
    {
      double a = __dadd_rn(0.0, __dmul_rn(sv[ps][0], dc[0][q]));
      a = __dadd_rn(a, __dmul_rn(sv[ps][1], dc[1][q]));
      ck[ps][q] = __dadd_rn(a, __dmul_rn(sv[ps][2], dc[2][q]));
    }
  }
  const double *zp = base + (size_t)3 * kTileSites * C;
  const unsigned int site0 = tile * kTileSites + lsite0;
  const unsigned int full =
      vec_end == ~0ull ? sites / 4 * 4 : (unsigned int)min((unsigned long long)sites, vec_end);
  if (C == 4)
  {
    double r[NQ];
    quad_transpose_sum<NQ>(ck, cat, r);
    const unsigned int lsite = lsite0 + quad_pass(cat) * kSitesPerPass;
    const double z = ld64_stream(zp + min(lsite, last));
    const unsigned int site = tile * kTileSites + lsite;
    if (site < sites)
    {
      r[0] = __dadd_rn(r[0], z);
      double terma = 0.0;
      if (WITH_LNL) { terma = __dmul_rn(r[0], m.rate_weight); if (!(terma > 0.0)) terma = terma != terma ? terma : 0.0; }
      site_terms<WITH_LNL>(m, r, terma, site, full, scv[0], out, lp);
    }
  }
  else
  {
#pragma unroll
    for (int ps = 0; ps < kPasses; ++ps)
    {
      const unsigned int site = site0 + ps * kSitesPerPass;
      double lk[NQ];
#pragma unroll
      for (int q = 0; q < NQ; ++q) lk[q] = ck[ps][q];
#pragma unroll
      for (int k = 1; k < C; ++k)
#pragma unroll
        for (int q = 0; q < NQ; ++q)
          lk[q] = __dadd_rn(lk[q], __shfl_down_sync(0xffffffffu, ck[ps][q], k, C));
      if (cat == 0 && site < sites)
      {
        lk[0] = __dadd_rn(lk[0], ld64_stream(zp + (lsite0 + ps * kSitesPerPass)));
        double terma = 0.0;
        if (WITH_LNL) { terma = __dmul_rn(lk[0], m.rate_weight); if (!(terma > 0.0)) terma = terma != terma ? terma : 0.0; }
        site_terms<WITH_LNL>(m, lk, terma, site, full, scv[ps], out, lp);
      }
    }
  }
}

// Compact sumtables, C = 4, one SITE per thread: the layout keeps a site's four categories next
// to each other, so a thread fetches a whole plane entry with one 32-byte load (3 x LDG.256 + one
// double of Z per site instead of 13 8-byte loads over four lanes), needs no shuffle for the
// category sums, and every lane carries a site through the division.  The 36 constants of the
// evaluation (4 categories x 3 live columns x {e, lr e, lr^2 e}) are read from shared memory
// (`cst`, broadcast).  Per-site arithmetic is that of sweep_tile_compact, term for term.
template <bool WITH_LNL>
__device__ __forceinline__ void sweep_tile_sites4(const ModelConst &m, const double *sumtable,
                                                  const unsigned int *sclr, unsigned int sites,
                                                  unsigned int tile, const double *cst, double out[3],
                                                  LnlProd &lp, unsigned long long vec_end)
{
  const double *base = sumtable + (size_t)tile * compact_tile_doubles<4>();
  const unsigned int last = sites - 1 - tile * kTileSites;
  const unsigned int ls = min((unsigned int)threadIdx.x, last);
  double sv[3][4];
#pragma unroll
  for (int p = 0; p < 3; ++p) ld256_stream(base + (size_t)p * kTileSites * 4 + (size_t)ls * 4, sv[p]);
  const double z = ld64_stream(base + (size_t)3 * kTileSites * 4 + ls);
  const unsigned int site = tile * kTileSites + threadIdx.x;
  unsigned int sc = 0u;
  if (WITH_LNL && sclr) sc = __ldg(sclr + tile * kTileSites + ls);
  double pair[2][3];
#pragma unroll
  for (int h = 0; h < 2; ++h)
  {
    double ck[2][3];
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
    {
      const int k = 2 * h + kk;
#pragma unroll
      for (int q = 0; q < 3; ++q)
      {
        double a = __dadd_rn(0.0, __dmul_rn(sv[0][k], cst[(k * 3 + 0) * 3 + q]));
        a = __dadd_rn(a, __dmul_rn(sv[1][k], cst[(k * 3 + 1) * 3 + q]));
        ck[kk][q] = __dadd_rn(a, __dmul_rn(sv[2][k], cst[(k * 3 + 2) * 3 + q]));
      }
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) pair[h][q] = __dadd_rn(ck[0][q], ck[1][q]);
  }
  double lk[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) lk[q] = __dadd_rn(pair[0][q], pair[1][q]);
  lk[0] = __dadd_rn(lk[0], z);
  if (site < sites)
  {
    const unsigned int full =
        vec_end == ~0ull ? sites / 4 * 4 : (unsigned int)min((unsigned long long)sites, vec_end);
    double terma = 0.0;
    if (WITH_LNL) { terma = __dmul_rn(lk[0], m.rate_weight); if (!(terma > 0.0)) terma = terma != terma ? terma : 0.0; }
    site_terms<WITH_LNL>(m, lk, terma, site, full, sc, out, lp);
  }
}

// Streaming form: the tiles [t0, t1) of one sumtable flow through a per-thread
// ring in shared memory.  Every thread copies ITS OWN entries of the tile that
// lies kRingStages-1 tiles ahead with cp.async (LDGSTS) and later reads them
// back, so the ring needs no barrier of any kind (thread-private slots, ordered
// by cp.async groups), and the bytes in flight no longer depend on registers:
// kRingStages-1 tiles x 32 KB per CTA are always on their way.
constexpr int kRingStages = 3;
#ifndef PT_SWEEP_RING
#define PT_SWEEP_RING 0   // measured on B200: plain LDG.256 streaming (3 CTAs/SM) beats the ring (2 CTAs/SM)
#endif

constexpr int kDefaultStages = PT_SWEEP_RING ? kRingStages : 0;   // 0 = direct loads

template <int C>
constexpr size_t ring_bytes(int stages) { return (size_t)stages * kPasses * kSitesPerPass * C * 32; }

template <int C>
constexpr size_t sweep_ring_bytes() { return ring_bytes<C>(kDefaultStages); }

// A launch with FEW work items, each a long chunk (one branch on a long alignment: B x nchunks
// of the order of the SM count, 30 tiles per chunk) cannot hide memory latency behind other
// CTAs -- there are none.  Such launches take the ring with as many stages as shared memory
// holds: 5 tiles (160 KB at C = 4) in flight per SM instead of one.  Same arithmetic in the
// same order as the direct form, so results are bit-identical.
template <int C>
constexpr int few_work_stages() { return C <= 4 ? 6 : 3; }

template <int C, bool WITH_LNL, int STAGES = kDefaultStages>
__device__ __forceinline__ void sweep_chunk(const ModelConst &m, const double *sumtable,
                                            const unsigned int *sclr, unsigned int sites,
                                            unsigned int t0, unsigned int t1, const double (&dg)[4][3],
                                            double *ring, double out[3], LnlProd &lp,
                                            unsigned long long vec_end, int compact = 0)
{
  if (compact > 0)
  {
    // compact sumtable (eigen column compact-1 folded into one double per site): direct loads
    for (unsigned int tile = t0; tile < t1; ++tile)
      sweep_tile_compact<C, WITH_LNL>(m, sumtable, sclr, sites, tile, dg, compact - 1, out, lp, vec_end);
    return;
  }
  if constexpr (STAGES == 0)
  {
    for (unsigned int tile = t0; tile < t1; ++tile)
      sweep_tile<C, WITH_LNL>(m, sumtable, sclr, sites, tile, dg, out, lp, vec_end);
  }
  else
  {
  constexpr int T = kSitesPerPass * C;
  const int cat = threadIdx.x % C;
  const unsigned int lsite = threadIdx.x / C;
  auto issue = [&](unsigned int tile, int stage) {
#pragma unroll
    for (int ps = 0; ps < kPasses; ++ps)
    {
      const unsigned int site = min(tile * kTileSites + lsite + ps * kSitesPerPass, sites - 1);
      const double *src = sumtable + ((size_t)site * C + cat) * 4;
      double *dst = ring + ((size_t)(stage * kPasses + ps) * T + threadIdx.x) * 4;
      cp_async16(dst, src);
      cp_async16(dst + 2, src + 2);
    }
  };
#pragma unroll
  for (int k = 0; k < STAGES - 1; ++k)
  {
    if (t0 + k < t1) issue(t0 + k, k);
    cp_async_commit();
  }
  int stage = 0, fill = STAGES - 1;
  for (unsigned int tile = t0; tile < t1; ++tile)
  {
    if (tile + (STAGES - 1) < t1) issue(tile + (STAGES - 1), fill);
    cp_async_commit();                       // (possibly empty) group: keeps the count uniform
    unsigned int scv[kPasses];
    load_scaler_words<C, WITH_LNL>(sclr, sites, tile, scv);
    cp_async_wait<STAGES - 1>();        // the group of `tile` has landed
    double sv[kPasses][4];
#pragma unroll
    for (int ps = 0; ps < kPasses; ++ps)
      lds256(ring + ((size_t)(stage * kPasses + ps) * T + threadIdx.x) * 4, sv[ps]);
    sweep_tile_compute<C, WITH_LNL>(m, sv, scv, sites, tile, dg, out, lp, vec_end);
    stage = stage + 1 == STAGES ? 0 : stage + 1;
    fill = fill + 1 == STAGES ? 0 : fill + 1;
  }
  cp_async_wait<0>();
  }
}

struct NewtonArgs
{
  ModelConst m;
  const double *sumtables;  // [B][entry_stride]
  size_t entry_stride;      // doubles between consecutive entries' sumtables
  const unsigned int *sclr; // [B][sites] summed scaler counts, or null (=> 0)
  NewtonState *st;          // [B]
  double *partials;         // [B][nchunks][3]: d1, d2 and (if want_lnl) lnL sums per chunk
  unsigned int sites, B, ntiles, chunk_tiles, nchunks, max_iter;
  // the cooperative kernel hands out `group` consecutive chunks of a branch per work item (a
  // scheduling unit only: partials stay per chunk, so results do not depend on it)
  unsigned int group;
  double eps, lambda;
  int want_lnl;
  int mode;                 // 0: Partition::OptimizeBranch, 1: pllmod_opt_minimize_newton
  double xmin, xmax;        // mode 1 bounds
  int compact;                // k + 1: sumtables are compact with eigen column k folded; 0: full sumtables
  double lnl_margin;          // predictor margin: carry the lnL when predicted |d1| < margin * eps
  unsigned long long vec_end; // ~0: unsharded; else local index where the global scalar tail starts
};

// One step of pllmod_opt_minimize_newton (PLLMOD/optimize/opt_algorithms.c:71-172)
// after the derivative evaluation number st.iters+1 at st.len: bracket update,
// Newton step with bisection fallback, stopping rules.  Plain double arithmetic
// without contraction, as the reference is compiled (-mavx, no FMA).
__device__ __forceinline__ void bounded_newton_step(NewtonState &st, double f, double df,
                                                    const NewtonArgs &a)
{
  const bool first = st.iters == 0;
  st.iters += 1;
  st.d1 = f;
  st.d2 = df;
  if (!isfinite(f) || !isfinite(df)) { st.status = 1; st.done = 1; return; }
  if (first)
  {
    if (df >= 0.0 && fabs(f) < a.eps) { st.done = 1; return; }
    if (f < 0.0) { st.xl = st.len; st.xh = a.xmax; }
    else { st.xh = st.len; st.xl = a.xmin; }
  }
  else
  {
    if (df > 0.0 && fabs(f) < a.eps) { st.done = 1; return; }
    if (f < 0.0) st.xl = st.len; else st.xh = st.len;
  }
  const double xl = st.xl, xh = st.xh;
  double rts = st.len, dx;
  const double rts_old = rts;
  const double t1 = __dadd_rn(__dmul_rn(__dadd_rn(rts, -xh), df), -f);
  const double t2 = __dadd_rn(__dmul_rn(__dadd_rn(rts, -xl), df), -f);
  if (df <= 0.0 || __dmul_rn(t1, t2) >= 0.0)
  {
    dx = __dmul_rn(0.5, __dadd_rn(xh, -xl));
    rts = __dadd_rn(xl, dx);
    if (xl == rts) { st.len = rts; st.done = 1; return; }
  }
  else
  {
    dx = __ddiv_rn(f, df);
    const double temp = rts;
    rts = __dadd_rn(rts, -dx);
    if (temp == rts) { st.len = rts; st.done = 1; return; }
  }
  if (fabs(dx) < a.eps) { st.len = rts_old; st.done = 1; return; }
  if ((unsigned int)st.iters >= a.max_iter) { st.len = rts_old; st.status = 2; st.done = 1; return; }
  if (rts < a.xmin) rts = a.xmin;
  st.len = rts;
}

// The per-branch scalar step after one derivative evaluation: mode 0 is the
// loop body of Partition::OptimizeBranch (PTPLL/pll_partition.cpp:433-481),
// mode 1 one step of pllmod_opt_minimize_newton.
__device__ __forceinline__ NewtonState newton_scalar_step(const NewtonArgs &a, const NewtonState *in,
                                                          double a1, double a2, double a3 = 0.0)
{
  NewtonState st;
  st.len = __ldcg(&in->len);
  const double evaluated_at = st.len;
  st.lnl = 0.0;
  st.lnl_valid = 0;
  const int had_lnl = __ldcg(&in->try_lnl);
  st.try_lnl = 0;
  st.lnl_sweeps = __ldcg(&in->lnl_sweeps) + (had_lnl ? 1 : 0);
  st.final_sweeps = 0;
  st.iters = __ldcg(&in->iters);
  st.done = 0;
  st.maybe_decreasing = __ldcg(&in->maybe_decreasing);
  st.status = 0;
  st.xl = __ldcg(&in->xl);
  st.xh = __ldcg(&in->xh);
  if (a.mode == 1)
    bounded_newton_step(st, a1, a2, a);
  else
  {
    double d1 = a1, d2 = a2;
    if (a.lambda != 0.0) d1 = __dadd_rn(d1, a.lambda);
    st.d1 = d1;
    st.d2 = d2;
    st.iters += 1;
    if (!isfinite(d1)) { st.status = 1; st.done = 1; }
    else if (!isfinite(d2)) { st.status = 2; st.done = 1; }
    else if (fabs(d1) < a.eps) st.done = 1;
    else
    {
      const double step = __ddiv_rn(d1, d2);
      st.len = (d2 < 0.0) ? __dadd_rn(st.len, step) : __dadd_rn(st.len, -step);
      // Will the next evaluation be the last?  Newton converges quadratically,
      // |d1(k+1)| ~ c |d1(k)|^2 with c estimated from the last two evaluations;
      // if that prediction is within 100x of the stopping threshold the next
      // sweep also sums the site log-likelihoods (a wrong guess costs either
      // one cheap extra log pass or one extra sweep at the end, never accuracy).
      {
        const double prev = __ldcg(&in->d1);
        const double pred = (st.iters > 1 && prev != 0.0)
                                ? fabs(d1) * (fabs(d1) / fabs(prev)) * (fabs(d1) / fabs(prev))
                                : fabs(d1);
        st.try_lnl = pred < a.lnl_margin * a.eps;
      }
      if (st.len < 0.0)
      {
        st.len = a.eps;
        if (st.maybe_decreasing) st.done = 1;
        st.maybe_decreasing = 1;
        st.try_lnl = 1;
      }
    }
    if (st.iters >= (int)a.max_iter) st.done = 1;
  }
  // the sweep that produced (a1, a2) also summed the site log-likelihoods at the
  // length it ran at; if that is where the branch ends up, the score is known
  if (a.want_lnl && had_lnl && st.done && st.len == evaluated_at)
  {
    st.lnl = a3;
    st.lnl_valid = 1;
  }
  return st;
}

__device__ __forceinline__ void newton_scalar_update(const NewtonArgs &a, unsigned int e, double a1,
                                                     double a2, double a3 = 0.0)
{
  a.st[e] = newton_scalar_step(a, &a.st[e], a1, a2, a3);
}

// Persistent cooperative kernel: the whole Newton-Raphson loop of
// Partition::OptimizeBranch (PTPLL/pll_partition.cpp:419-481) for B branches
// at once.  Each iteration = one sweep of every still-running branch's
// sumtable (chunks of tiles spread over all CTAs), a grid barrier, a
// fixed-order reduction + the scalar update done by one CTA per branch, a grid
// barrier.  Reduction units (chunks) depend only on the site count, so results
// are independent of the grid size.
template <int C, bool LNL, int STAGES = kDefaultStages>
__global__ void __launch_bounds__(kSitesPerPass *C, (C <= 4 && STAGES <= 3 ? (STAGES ? 2 : 3) : 1)) newton_kernel(NewtonArgs a)
{
  constexpr int T = kSitesPerPass * C;
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) double ring[];
  __shared__ double red[3 * (T / 32 < 1 ? 1 : T / 32)];
  // consecutive chunks per work item: re-chosen every iteration from the number of branches still
  // running, so that the last iterations of a batch (a handful of slow branches) still spread over
  // the whole grid.  A scheduling unit only: partials stay per chunk.
  const unsigned int max_group = a.group ? a.group : 1u;
  unsigned int group = max_group, ngroups = (a.nchunks + group - 1) / group, nwork = a.B * ngroups;
  __shared__ unsigned int sh_running;
  const int cat = threadIdx.x % C;
  // Compact sumtables with four categories take the site-per-thread sweep; its chunk partials are
  // PER WARP (written straight after the warp's shuffle tree: no CTA barrier anywhere in a sweep,
  // so a CTA's loads never drain between chunks), the reduction below sums nchunks x W of them.
  constexpr int W = T / 32 < 1 ? 1 : T / 32;
  const bool sites4 = C == 4 && a.compact > 0;
  const unsigned int nparts = sites4 ? a.nchunks * W : a.nchunks;
  __shared__ double cst[36];

  auto sweep_item_sites4 = [&](unsigned int e, unsigned int c0, unsigned int c1, bool final_pass) {
    __syncthreads();   // the previous work item's constants are no longer read
    if (threadIdx.x < 4)
    {
      double dg[4][3];
      diag_regs<C>(a.m, __ldcg(&a.st[e].len), threadIdx.x, dg);
#pragma unroll
      for (int p = 0; p < 3; ++p)
      {
        const int j = p + (p >= a.compact - 1 ? 1 : 0);
#pragma unroll
        for (int q = 0; q < 3; ++q)
          cst[(threadIdx.x * 3 + p) * 3 + q] = j == 1 ? dg[1][q] : (j == 2 ? dg[2][q] : (j == 3 ? dg[3][q] : dg[0][q]));
      }
    }
    __syncthreads();
    const bool with_lnl = LNL && (final_pass || __ldcg(&a.st[e].try_lnl));
    const double *table = a.sumtables + (size_t)e * a.entry_stride;
    const unsigned int *sclr = a.sclr ? a.sclr + (size_t)e * a.sites : nullptr;
    for (unsigned int chunk = c0; chunk < c1; ++chunk)
    {
      double out[3] = {0.0, 0.0, 0.0};
      LnlProd lp;
      const unsigned int t1 = min((chunk + 1) * a.chunk_tiles, a.ntiles);
      if (with_lnl)
        for (unsigned int tile = chunk * a.chunk_tiles; tile < t1; ++tile)
          sweep_tile_sites4<true>(a.m, table, sclr, a.sites, tile, cst, out, lp, a.vec_end);
      else
        for (unsigned int tile = chunk * a.chunk_tiles; tile < t1; ++tile)
          sweep_tile_sites4<false>(a.m, table, nullptr, a.sites, tile, cst, out, lp, a.vec_end);
      double v[3] = {out[0], out[1], (LNL && with_lnl) ? lp.value() : 0.0};
#pragma unroll
      for (int k = 0; k < (LNL ? 3 : 2); ++k)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v[k] = __dadd_rn(v[k], __shfl_down_sync(0xffffffffu, v[k], d));
      if ((threadIdx.x & 31) == 0)
      {
        double *dst = a.partials + (((size_t)e * a.nchunks + chunk) * W + (threadIdx.x >> 5)) * 3;
        if (!final_pass) { dst[0] = v[0]; dst[1] = v[1]; }
        if (LNL) dst[2] = v[2];
      }
    }
  };

  // one sweep of the chunks [c0, c1) of branch e at its current length; FINAL: lnL only
  auto sweep_item = [&](unsigned int e, unsigned int c0, unsigned int c1, bool final_pass) {
    if (sites4) { sweep_item_sites4(e, c0, c1, final_pass); return; }
    double dg[4][3];
    diag_regs<C>(a.m, __ldcg(&a.st[e].len), cat, dg);
    const bool with_lnl = LNL && (final_pass || __ldcg(&a.st[e].try_lnl));
    for (unsigned int chunk = c0; chunk < c1; ++chunk)
    {
      double out[3] = {0.0, 0.0, 0.0};
      LnlProd lp;
      const unsigned int t1 = min((chunk + 1) * a.chunk_tiles, a.ntiles);
      if (with_lnl)
        sweep_chunk<C, true, STAGES>(a.m, a.sumtables + (size_t)e * a.entry_stride,
                              a.sclr ? a.sclr + (size_t)e * a.sites : nullptr, a.sites,
                              chunk * a.chunk_tiles, t1, dg, ring, out, lp, a.vec_end, a.compact);
      else
        sweep_chunk<C, false, STAGES>(a.m, a.sumtables + (size_t)e * a.entry_stride, nullptr, a.sites,
                               chunk * a.chunk_tiles, t1, dg, ring, out, lp, a.vec_end, a.compact);
      double *dst = a.partials + ((size_t)e * a.nchunks + chunk) * 3;
      if (LNL)
      {
        double s[3] = {out[0], out[1], with_lnl ? lp.value() : 0.0};
        block_sum_n<T, 3>(s, red);
        if (threadIdx.x == 0)
        {
          if (!final_pass) { dst[0] = s[0]; dst[1] = s[1]; }
          dst[2] = s[2];
        }
      }
      else
      {
        double s[2] = {out[0], out[1]};
        block_sum_n<T, 2>(s, red);
        if (threadIdx.x == 0) { dst[0] = s[0]; dst[1] = s[1]; }
      }
    }
  };

  // One loop for the Newton iterations and the (rare) final lnL-only pass, so that the sweep
  // code has a single call site and inlines (the arguments stay in the parameter bank).
  bool final_pass = false;
  for (unsigned int it = 0;; ++it)
  {
    {
      if (threadIdx.x == 0) sh_running = 0u;
      __syncthreads();
      unsigned int mine = 0u;
      for (unsigned int e = threadIdx.x; e < a.B; e += T)
        mine += final_pass ? !__ldcg(&a.st[e].lnl_valid) : !__ldcg(&a.st[e].done);
      if (mine) atomicAdd(&sh_running, mine);
      __syncthreads();
      const unsigned int running = sh_running;   // the same in every CTA: flags are stable here
      group = max(1u, min(max_group, running * a.nchunks / (8u * gridDim.x)));
      ngroups = (a.nchunks + group - 1) / group;
      nwork = a.B * ngroups;
    }
    for (unsigned int w = blockIdx.x; w < nwork; w += gridDim.x)
    {
      const unsigned int e = w / ngroups, g = w % ngroups;
      if (final_pass ? __ldcg(&a.st[e].lnl_valid) : __ldcg(&a.st[e].done)) continue;
      sweep_item(e, g * group, min((g + 1) * group, a.nchunks), final_pass);
    }
    grid.sync();
    for (unsigned int e = blockIdx.x; e < a.B; e += gridDim.x)
    {
      if (final_pass ? __ldcg(&a.st[e].lnl_valid) : __ldcg(&a.st[e].done)) continue;
      double s[3] = {0.0, 0.0, 0.0};
      for (unsigned int t = threadIdx.x; t < nparts; t += T)
      {
        if (!final_pass)
        {
          s[0] = __dadd_rn(s[0], __ldcg(a.partials + ((size_t)e * nparts + t) * 3 + 0));
          s[1] = __dadd_rn(s[1], __ldcg(a.partials + ((size_t)e * nparts + t) * 3 + 1));
        }
        if (LNL) s[2] = __dadd_rn(s[2], __ldcg(a.partials + ((size_t)e * nparts + t) * 3 + 2));
      }
      block_sum_n<T, 3>(s, red);
      if (threadIdx.x == 0)
      {
        if (!final_pass)
          newton_scalar_update(a, e, s[0], s[1], s[2]);
        else
        {
          a.st[e].lnl = s[2];
          a.st[e].lnl_valid = 1;
          a.st[e].final_sweeps = 1;
        }
      }
    }
    if (final_pass) break;
    grid.sync();
    int active = 0;
    for (unsigned int e = threadIdx.x; e < a.B; e += T) active |= !__ldcg(&a.st[e].done);
    if (__syncthreads_or(active) && it + 1 < a.max_iter) continue;
    if (!LNL) break;
    // branches that stopped somewhere else than at the last evaluated length (second negative
    // step, iteration limit): one more sweep at the final length, for the lnL only
    int pending = 0;
    for (unsigned int e = threadIdx.x; e < a.B; e += T) pending |= !__ldcg(&a.st[e].lnl_valid);
    if (!__syncthreads_or(pending)) break;   // uniform over the grid: every CTA reads the same flags
    final_pass = true;
  }
}

// ---- split-phase Newton for site shards -----------------------
// The same three phases as the cooperative kernel, as separate launches so that
// the caller's sum over shards (the allreduce hook, one [2B] or [B] buffer per
// iteration) can run between the local reduction and the scalar update -- the
// scheme of pll-modules' parallel_reduce_cb (PLLMOD/optimize/pll_optimize.c:1211-1255).
template <int C, bool LNL, bool FINAL, int STAGES = kDefaultStages>
__global__ void __launch_bounds__(kSitesPerPass *C) newton_sweep_kernel(NewtonArgs a)
{
  constexpr int T = kSitesPerPass * C;
  extern __shared__ __align__(16) double ring[];
  __shared__ double red[T / 32 < 1 ? 1 : T / 32];
  const unsigned int nwork = a.B * a.nchunks;
  const int cat = threadIdx.x % C;
  for (unsigned int w = blockIdx.x; w < nwork; w += gridDim.x)
  {
    const unsigned int e = w / a.nchunks, chunk = w % a.nchunks;
    if (FINAL ? a.st[e].lnl_valid : a.st[e].done) continue;
    double dg[4][3];
    diag_regs<C>(a.m, a.st[e].len, cat, dg);
    double out[3] = {0.0, 0.0, 0.0};
    LnlProd lp;
    const unsigned int t1 = min((chunk + 1) * a.chunk_tiles, a.ntiles);
    const bool with_lnl = LNL && (FINAL || a.st[e].try_lnl);
    if (with_lnl)
      sweep_chunk<C, true, STAGES>(a.m, a.sumtables + (size_t)e * a.entry_stride,
                              a.sclr ? a.sclr + (size_t)e * a.sites : nullptr, a.sites,
                              chunk * a.chunk_tiles, t1, dg, ring, out, lp, a.vec_end, a.compact);
    else
      sweep_chunk<C, false, STAGES>(a.m, a.sumtables + (size_t)e * a.entry_stride, nullptr, a.sites,
                               chunk * a.chunk_tiles, t1, dg, ring, out, lp, a.vec_end, a.compact);
    const double s1 = block_sum<T>(out[0], red);
    const double s2 = block_sum<T>(out[1], red);
    const double s3 = LNL ? block_sum<T>(with_lnl ? lp.value() : 0.0, red) : 0.0;
    if (threadIdx.x == 0)
    {
      a.partials[(size_t)w * 3 + 0] = s1;
      a.partials[(size_t)w * 3 + 1] = s2;
      a.partials[(size_t)w * 3 + 2] = s3;
    }
  }
}

// one CTA per branch: chunk partials -> shard[e*3 .. e*3+2] in fixed order (zeros
// for idle branches, so the cross-shard sum stays well defined)
template <bool FINAL>
__global__ void __launch_bounds__(256) newton_reduce_kernel(NewtonArgs a, double *shard)
{
  __shared__ double red[8];
  const unsigned int e = blockIdx.x;
  const bool idle = FINAL ? a.st[e].lnl_valid : a.st[e].done;
#pragma unroll
  for (int v = 0; v < 3; ++v)
  {
    double acc = 0.0;
    if (!idle)
      for (unsigned int t = threadIdx.x; t < a.nchunks; t += 256)
        acc = __dadd_rn(acc, a.partials[((size_t)e * a.nchunks + t) * 3 + v]);
    acc = block_sum<256>(acc, red);
    if (threadIdx.x == 0) shard[(size_t)e * 3 + v] = acc;
  }
}

// scalar updates from the (summed) shard buffer; *active receives the number of
// branches still running, active[1] the number still lacking their final lnL
__global__ void newton_update_kernel(NewtonArgs a, const double *shard, unsigned int *active)
{
  const unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= a.B) return;
  if (!a.st[e].done)
    newton_scalar_update(a, e, shard[(size_t)e * 3], shard[(size_t)e * 3 + 1], shard[(size_t)e * 3 + 2]);
  if (!a.st[e].done) atomicAdd(active, 1u);
  else if (a.want_lnl && !a.st[e].lnl_valid) atomicAdd(active + 1, 1u);
}

__global__ void newton_store_lnl_kernel(NewtonArgs a, const double *shard)
{
  const unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < a.B && !a.st[e].lnl_valid)
  {
    a.st[e].lnl = shard[(size_t)e * 3 + 2];
    a.st[e].lnl_valid = 1;
  }
}

// Sum of `count` doubles at `buf` over all site shards, in place, on the engine's stream: the
// exchange of the single-purpose entry points (one launch of one CTA; the persistent
// step-program kernel does the same exchange inline).  count <= kShardSlots.
__global__ void __launch_bounds__(256) shard_allreduce_kernel(ShardComm c, double *buf, unsigned int count)
{
  const unsigned long long q = *(volatile unsigned long long *)c.seq;
  for (unsigned int i = threadIdx.x; i < count; i += blockDim.x) shard_put(c, q, i, buf[i]);
  for (unsigned int i = threadIdx.x; i < count; i += blockDim.x)
  {
    double acc = shard_get(c, q, i, 0);
    for (int r = 1; r < c.nranks; ++r) acc = __dadd_rn(acc, shard_get(c, q, i, r));
    buf[i] = acc;
  }
  __syncthreads();   // everybody has read the sequence number and collected its values
  if (threadIdx.x == 0) *c.seq = q + 1;
}

// edge log-likelihood of a shard: tile partials -> one double in device memory
__global__ void __launch_bounds__(256) sum_partials_kernel(const double *partials, unsigned int n,
                                                           double *out)
{
  __shared__ double red[8];
  double acc = 0.0;
  for (unsigned int t = threadIdx.x; t < n; t += 256) acc = __dadd_rn(acc, partials[t]);
  acc = block_sum<256>(acc, red);
  if (threadIdx.x == 0) *out = acc;
}

// ------------------------------------------------- NNI candidate preparation
// For one candidate and one tile of sites: the two CLV updates that face each
// other across the swapped edge (kept in registers, never written), their
// per-site scaler counts, and the sumtable entries of the central edge.
// Reads 4 neighbour CLVs (or tip masks) + their scaler words; writes the
// candidate's sumtable (32 B per entry) and one summed scaler word per site.
// The candidate index is the fastest-varying part of blockIdx.x, so the CTAs
// that run together work on the SAME tile of sites for different candidates:
// each directional CLV tile is fetched from HBM once and then served from L2
// to the ~2.6 candidates that use it.
// Shared-memory staging of one candidate's four neighbour tiles.  Each child is
// a contiguous run in HBM (CLV tile kHeavyTile*C*32 B, scaler tile, or tip-mask
// tile), fetched by ONE thread with TMA bulk copies that complete on the
// child's mbarrier; all 4 x (16 KB + 0.5 KB) are in flight at once per CTA, so
// memory-level parallelism no longer depends on registers or warp count.
template <int C>
struct PrepSmem
{
  double clv[4][kHeavyTile * C * 4];
  unsigned int scal[4][kHeavyTile];
  unsigned char tip[4][kHeavyTile];
  unsigned long long bar[4];
};

template <int C>
__device__ __forceinline__ void stage_child(const DevState &s, PrepSmem<C> &sm, int slot,
                                            unsigned int clv_index, int scaler_index,
                                            unsigned int site_base, unsigned int nvalid)
{
  if (clv_index < s.n_tips)
  {
    const unsigned int bytes = (nvalid + 15u) & ~15u;
    mbar_expect_tx(&sm.bar[slot], bytes);
    bulk_g2s(sm.tip[slot], s.tips + (size_t)clv_index * s.tip_stride + site_base, bytes,
             &sm.bar[slot]);
    return;
  }
  const unsigned int cbytes = nvalid * C * 32u;
  const unsigned int sbytes = scaler_index >= 0 ? ((nvalid * 4u + 15u) & ~15u) : 0u;
  mbar_expect_tx(&sm.bar[slot], cbytes + sbytes);
  bulk_g2s(sm.clv[slot],
           s.clv + (size_t)(clv_index - s.n_tips) * s.clv_stride + (size_t)site_base * C * 4,
           cbytes, &sm.bar[slot]);
  if (sbytes)
    bulk_g2s(sm.scal[slot], s.scalers + (size_t)scaler_index * s.scaler_stride + site_base,
             sbytes, &sm.bar[slot]);
}

template <int C>
__device__ __forceinline__ unsigned int staged_entry(const DevState &s, const PrepSmem<C> &sm,
                                                     int slot, unsigned int clv_index,
                                                     int scaler_index, unsigned int local_site,
                                                     int cat, double v[4])
{
  if (clv_index < s.n_tips)
  {
    const unsigned int m = sm.tip[slot][local_site];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ((m >> j) & 1u) ? 1.0 : 0.0;
    return 0u;
  }
  lds256(&sm.clv[slot][((size_t)local_site * C + cat) * 4], v);
  return scaler_index >= 0 ? sm.scal[slot][local_site] : 0u;
}

// One CTA = one candidate group x one tile of sites.  The four projected
// vectors M_k * child_k (the expensive part: 4 dot products each) are computed
// once and shared by the two candidates of the group; per candidate only the
// elementwise products, the scaling test and the sumtable entry remain.
template <int C>
__device__ __forceinline__ void finish_candidate(const DevState &s, const ModelConst &m,
                                                 const double a[4], const double b[4],
                                                 const double c[4], const double d[4],
                                                 unsigned int cnt, bool store, double *table,
                                                 unsigned int site, unsigned int *sclr_dst, int cat,
                                                 int compact = -1)
{
  double pp[4], qq[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    pp[i] = __dmul_rn(a[i], b[i]);
    qq[i] = __dmul_rn(c[i], d[i]);
  }
  bool sp = (pp[0] < PT_SCALE_THRESHOLD) & (pp[1] < PT_SCALE_THRESHOLD) &
            (pp[2] < PT_SCALE_THRESHOLD) & (pp[3] < PT_SCALE_THRESHOLD);
  bool sq = (qq[0] < PT_SCALE_THRESHOLD) & (qq[1] < PT_SCALE_THRESHOLD) &
            (qq[2] < PT_SCALE_THRESHOLD) & (qq[3] < PT_SCALE_THRESHOLD);
  sp = site_all<C>(sp);
  sq = site_all<C>(sq);
  if (sp)
  {
#pragma unroll
    for (int i = 0; i < 4; ++i) pp[i] = __dmul_rn(pp[i], PT_SCALE_FACTOR);
  }
  if (sq)
  {
#pragma unroll
    for (int i = 0; i < 4; ++i) qq[i] = __dmul_rn(qq[i], PT_SCALE_FACTOR);
  }
  double sum[4];
  sumtable_entry<C>(m, pp, qq, sum);
  if (compact >= 0)
  {
    // compact form (see sweep_tile_compact): the three live columns in their planes, and the
    // folded column summed over the site's categories (c0 + c1 + ... in order) by lane cat 0
    const double zc = compact == 0 ? sum[0] : (compact == 1 ? sum[1] : (compact == 2 ? sum[2] : sum[3]));
    double z = zc;
#pragma unroll
    for (int k = 1; k < C; ++k) z = __dadd_rn(z, __shfl_down_sync(0xffffffffu, zc, k, C));
    if (store)
    {
      double *base = table + (size_t)(site / kTileSites) * compact_tile_doubles<C>();
      const unsigned int ls = site % kTileSites;
#pragma unroll
      for (int p = 0; p < 3; ++p)
      {
        const int j = p + (p >= compact ? 1 : 0);
        base[(size_t)p * kTileSites * C + (size_t)ls * C + cat] =
            j == 1 ? sum[1] : (j == 2 ? sum[2] : (j == 3 ? sum[3] : sum[0]));
      }
      if (cat == 0)
      {
        base[(size_t)3 * kTileSites * C + ls] = z;
        *sclr_dst = cnt + (sp ? 1u : 0u) + (sq ? 1u : 0u);
      }
    }
    return;
  }
  if (store)
  {
    st256(table + ((size_t)site * C + cat) * 4, sum);
    if (cat == 0) *sclr_dst = cnt + (sp ? 1u : 0u) + (sq ? 1u : 0u);
  }
}

template <int C>
__global__ void __launch_bounds__(kHeavySites *C, PT_PREP_MINB)
    nni_prepare_kernel(DevState s, ModelConst m, const DevGroup *__restrict__ groups,
                       unsigned int ngroups, double *sumtables, size_t entry_stride,
                       unsigned int *sclr)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  PrepSmem<C> &sm = *reinterpret_cast<PrepSmem<C> *>(smem_raw);
  const unsigned int g = blockIdx.x % ngroups, tile = blockIdx.x / ngroups;
  const DevGroup gr = groups[g];
  const int cat = threadIdx.x % C;
  const unsigned int lsite0 = threadIdx.x / C;
  const unsigned int site_base = tile * kHeavyTile;
  const unsigned int nvalid = min((unsigned int)kHeavyTile, s.sites - site_base);

  if (threadIdx.x == 0)
  {
#pragma unroll
    for (int i = 0; i < 4; ++i) mbar_init(&sm.bar[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0)
  {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      stage_child<C>(s, sm, k, gr.ch[k].clv, gr.ch[k].scaler, site_base, nvalid);
  }

  double proj[4][kHeavyPasses][4];
  unsigned int cnt[kHeavyPasses];
#pragma unroll
  for (int ps = 0; ps < kHeavyPasses; ++ps) cnt[ps] = 0u;
#pragma unroll
  for (int k = 0; k < 4; ++k)
  {
    double mt[16];
    load_matrix<C>(s.pmat + ((size_t)gr.ch[k].mat * C + cat) * 16, mt);
    mbar_wait(&sm.bar[k], 0);
#pragma unroll
    for (int ps = 0; ps < kHeavyPasses; ++ps)
    {
      const unsigned int ls = min(lsite0 + ps * kHeavySites, nvalid - 1);
      double x[4];
      cnt[ps] += staged_entry<C>(s, sm, k, gr.ch[k].clv, gr.ch[k].scaler, ls, cat, x);
#pragma unroll
      for (int i = 0; i < 4; ++i) proj[k][ps][i] = dot4(mt + 4 * i, x);
    }
  }
#pragma unroll
  for (int ps = 0; ps < kHeavyPasses; ++ps)
  {
    const unsigned int ls_raw = lsite0 + ps * kHeavySites;
    const bool ok = ls_raw < nvalid;
    const unsigned int site = site_base + min(ls_raw, nvalid - 1);
    finish_candidate<C>(s, m, proj[0][ps], proj[1][ps], proj[2][ps], proj[3][ps], cnt[ps], ok,
                        sumtables + (size_t)gr.out[0] * entry_stride, site,
                        sclr + (size_t)gr.out[0] * s.sites + site, cat);
    if (gr.ncand > 1)
      finish_candidate<C>(s, m, proj[3][ps], proj[1][ps], proj[2][ps], proj[0][ps], cnt[ps], ok,
                          sumtables + (size_t)gr.out[1] * entry_stride, site,
                          sclr + (size_t)gr.out[1] * s.sites + site, cat);
  }
}

// ---- NNI prepare, version 2 ---------------------------------------------------
// Same staging (TMA bulk copies on four mbarriers) and same arithmetic as
// nni_prepare_kernel, with the register footprint cut so that three CTAs fit an
// SM: the four P matrices are staged once per CTA in shared memory (padded
// category blocks, broadcast LDS.128 rows) instead of 32 registers per thread,
// and the tile is walked two passes at a time (64 projection registers instead
// of 128).
template <int C>
struct PrepSmem2
{
  double clv[4][kHeavyTile * C * 4];
  unsigned int scal[4][kHeavyTile];
  unsigned char tip[4][kHeavyTile];
  unsigned long long bar[4];
  // everything above has the layout of PrepSmem<C> (stage_child / staged_entry use that view)
  __align__(16) double mat[4][C * kMatPad];
};

template <int C>
__global__ void __launch_bounds__(kHeavySites *C, (C <= 4 ? 3 : 1))
    nni_prepare_kernel_v2(DevState s, ModelConst m, const DevGroup *__restrict__ groups,
                          unsigned int ngroups, double *sumtables, size_t entry_stride,
                          unsigned int *sclr, int compact)
{
  extern __shared__ __align__(128) unsigned char smem_raw[];
  PrepSmem2<C> &sm = *reinterpret_cast<PrepSmem2<C> *>(smem_raw);
  const unsigned int g = blockIdx.x % ngroups, tile = blockIdx.x / ngroups;
  const DevGroup gr = groups[g];
  const int cat = threadIdx.x % C;
  const unsigned int lsite0 = threadIdx.x / C;
  const unsigned int site_base = tile * kHeavyTile;
  const unsigned int nvalid = min((unsigned int)kHeavyTile, s.sites - site_base);

  // thread 0 arms the barriers and starts the four bulk copies at once (same
  // thread: program order); everybody else stages the matrices meanwhile, and
  // one CTA barrier publishes both the barriers and the matrices
  if (threadIdx.x == 0)
  {
#pragma unroll
    for (int i = 0; i < 4; ++i) mbar_init(&sm.bar[i], 1);
    mbar_fence_init();
#pragma unroll
    for (int k = 0; k < 4; ++k)
      stage_child<C>(s, *reinterpret_cast<PrepSmem<C> *>(smem_raw), k, gr.ch[k].clv, gr.ch[k].scaler,
                     site_base, nvalid);
  }
  // the four matrices: 4 x C x 16 doubles, one double2 per thread and step
  for (int c = threadIdx.x; c < 4 * C * 8; c += kHeavySites * C)
  {
    const int k = c / (C * 8), rest = c % (C * 8), mc = rest / 8, within = rest % 8;
    const double2 v = *reinterpret_cast<const double2 *>(
        s.pmat + ((size_t)gr.ch[k].mat * C + mc) * 16 + within * 2);
    *reinterpret_cast<double2 *>(&sm.mat[k][mc * kMatPad + within * 2]) = v;
  }
  __syncthreads();

#pragma unroll 1
  for (int half = 0; half < kHeavyPasses; half += 2)
  {
    double proj[4][2][4];
    unsigned int cnt[2] = {0u, 0u};
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
      if (half == 0) mbar_wait(&sm.bar[k], 0);
      double x[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int ls = min(lsite0 + (half + j) * kHeavySites, nvalid - 1);
        cnt[j] += staged_entry<C>(s, *reinterpret_cast<PrepSmem<C> *>(smem_raw), k, gr.ch[k].clv,
                                  gr.ch[k].scaler, ls, cat, x[j]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
      {
        double row[4];
        lds256(&sm.mat[k][cat * kMatPad + 4 * i], row);
#pragma unroll
        for (int j = 0; j < 2; ++j) proj[k][j][i] = dot4(row, x[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
    {
      const unsigned int ls_raw = lsite0 + (half + j) * kHeavySites;
      const bool ok = ls_raw < nvalid;
      const unsigned int site = site_base + min(ls_raw, nvalid - 1);
      finish_candidate<C>(s, m, proj[0][j], proj[1][j], proj[2][j], proj[3][j], cnt[j], ok,
                          sumtables + (size_t)gr.out[0] * entry_stride, site,
                          sclr + (size_t)gr.out[0] * s.sites + site, cat, compact - 1);
      if (gr.ncand > 1)
        finish_candidate<C>(s, m, proj[3][j], proj[1][j], proj[2][j], proj[0][j], cnt[j], ok,
                            sumtables + (size_t)gr.out[1] * entry_stride, site,
                            sclr + (size_t)gr.out[1] * s.sites + site, cat, compact - 1);
    }
  }
}

} // namespace ptgpu
