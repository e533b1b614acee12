/*
 * oracle/refglue/ref_capi.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A flat C API over the UNMODIFIED reference (pt::pll::Partition from
 * /root/reference/lib/libptpll/src + libpll + pll-modules, compiled in place
 * by oracle/Makefile.ref) so that Python (ctypes) can drive it: golden-vector
 * generation (tests/golden/make_golden.py), oracle validation
 * (tests/test_oracle.py) and the timed CPU baseline of bench.py
 * (cpu_baseline.kind == "reference").  Lives only in oracle/_ref/libpllref.so.
 *
 * Nodes are addressed by libpll's node_index (unique per pll_unode_t record:
 * tips 0..n-1, inner records n..n+3(n-2)-1; LIBPLL/parse_utree.y:250-340).
 */
#include <cstring>
#include <map>
#include <string>
#include <vector>

#define private public /* test glue reads Partition::partition_ directly */
#include <pll_partition.hpp>
#undef private
#include <model.hpp>
#include <pll_util.hpp>

extern "C" {
#include <libpll/pll_optimize.h>
}

using pt::pll::Model;
using pt::pll::Partition;
using pt::pll::TraversalType;

namespace {

struct RefHandle {
  std::vector<pll_utree_t*> trees;
  pll_utree_t* tree = nullptr;
  Partition* part = nullptr;
  Model model;
  std::vector<pll_unode_t*> by_index;
  std::string error;
};

thread_local std::string g_error;

void index_nodes(RefHandle* h)
{
  pll_utree_t* t = h->tree;
  h->by_index.assign(t->tip_count + 3 * t->inner_count, nullptr);
  for (unsigned i = 0; i < t->tip_count + t->inner_count; ++i) {
    pll_unode_t* n = t->nodes[i];
    h->by_index.at(n->node_index) = n;
    if (n->next) {
      h->by_index.at(n->next->node_index) = n->next;
      h->by_index.at(n->next->next->node_index) = n->next->next;
    }
  }
}

} // namespace

extern "C" {

const char* ref_last_error() { return g_error.c_str(); }

/* newick may hold several trees (one per line); tree_index picks one.
   alpha <= 0: take everything from the RAxML_info file. */
void* ref_open(const char* newick_path, const char* fasta_path,
               const char* raxml_info_path, int rate_cats, int tree_index,
               int map_mode)
{
  try {
    RefHandle* h = new RefHandle;
    h->trees = pt::pll::ParseMultiNewick(newick_path);
    h->tree = h->trees.at(tree_index);
    std::vector<std::string> labels, seqs;
    pt::pll::ParseFasta(fasta_path, h->tree->tip_count, labels, seqs);
    h->model = pt::pll::ParseRaxmlInfo(raxml_info_path, rate_cats);
    h->part = new Partition(h->tree, h->model, labels, seqs, map_mode != 0);
    index_nodes(h);
    return h;
  } catch (const std::exception& e) {
    g_error = e.what();
    return nullptr;
  }
}

/* same, but from in-memory strings and an explicit model (synthetic data) */
void* ref_open_mem(const char* newick, int tips, const char** labels,
                   const char** seqs, const double* freqs, const double* subst,
                   const double* rates, int rate_cats, int map_mode)
{
  try {
    RefHandle* h = new RefHandle;
    pll_utree_t* t = pll_utree_parse_newick_string(newick);
    if (!t) throw std::runtime_error("newick parse failed");
    h->trees.push_back(t);
    h->tree = t;
    std::vector<std::string> l(labels, labels + tips), s(seqs, seqs + tips);
    h->model = Model{"GTR", std::vector<double>(freqs, freqs + 4),
                     std::vector<double>(subst, subst + 6),
                     std::vector<double>(rates, rates + rate_cats)};
    h->part = new Partition(t, h->model, l, s, map_mode != 0);
    index_nodes(h);
    return h;
  } catch (const std::exception& e) {
    g_error = e.what();
    return nullptr;
  }
}

/* the reference's own pattern compression and weights (never called by pt; used to pin the
   engine's pattern-weight path): sequences are rewritten in place, returns the new length */
int ref_compress(char** seqs, int count, int length, unsigned* weights_out)
{
  int len = length;
  unsigned int* w = pll_compress_site_patterns(seqs, pll_map_nt, count, &len);
  if (!w) return -1;
  memcpy(weights_out, w, sizeof(unsigned) * len);
  free(w);
  return len;
}

void ref_set_pattern_weights(void* hv, const unsigned* weights)
{
  pll_set_pattern_weights(static_cast<RefHandle*>(hv)->part->partition_.get(), weights);
}

void ref_close(void* hv)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  delete h->part;
  for (auto t : h->trees) pll_utree_destroy(t, pt::pll::cb_erase_data);
  delete h;
}

/* switch to another tree of the same multi-Newick file */
int ref_select_tree(void* hv, int tree_index)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  if (tree_index < 0 || (size_t)tree_index >= h->trees.size()) return 0;
  h->tree = h->trees[tree_index];
  index_nodes(h);
  return 1;
}

unsigned ref_tips(void* hv) { return static_cast<RefHandle*>(hv)->tree->tip_count; }
unsigned ref_sites(void* hv) { return static_cast<RefHandle*>(hv)->part->partition_->sites; }
unsigned ref_cats(void* hv) { return static_cast<RefHandle*>(hv)->part->partition_->rate_cats; }
unsigned ref_record_count(void* hv) { return static_cast<RefHandle*>(hv)->by_index.size(); }
unsigned ref_root_index(void* hv)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  return pt::pll::GetVirtualRoot(h->tree)->node_index;
}

/* tree->nodes[] order (tips then inner, LIBPLL/parse_utree.y:395-445) */
void ref_nodes_order(void* hv, unsigned* out)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  for (unsigned i = 0; i < h->tree->tip_count + h->tree->inner_count; ++i)
    out[i] = h->tree->nodes[i]->node_index;
}

/* per record: next, back (node_index or -1), clv, scaler, pmatrix */
void ref_record(void* hv, unsigned idx, int* next, int* back, unsigned* clv,
                int* scaler, unsigned* pmatrix, double* length,
                char* label, unsigned label_cap)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  pll_unode_t* n = h->by_index.at(idx);
  *next = n->next ? (int)n->next->node_index : -1;
  *back = n->back ? (int)n->back->node_index : -1;
  *clv = n->clv_index;
  *scaler = n->scaler_index;
  *pmatrix = n->pmatrix_index;
  *length = n->length;
  if (label && label_cap) {
    label[0] = 0;
    if (n->label) { strncpy(label, n->label, label_cap - 1); label[label_cap - 1] = 0; }
  }
}

void ref_model(void* hv, double* freqs, double* subst, double* rates)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  Model m = h->part->GetModel();
  memcpy(freqs, m.frequencies.data(), 4 * sizeof(double));
  memcpy(subst, m.subst_params.data(), 6 * sizeof(double));
  memcpy(rates, m.category_rates.data(), m.category_rates.size() * sizeof(double));
}

void ref_eigen(void* hv, double* evecs, double* inv_evecs, double* evals)
{
  pll_partition_t* p = static_cast<RefHandle*>(hv)->part->partition_.get();
  if (!p->eigen_decomp_valid[0]) pll_update_eigen(p, 0);
  memcpy(evecs, p->eigenvecs[0], 16 * sizeof(double));
  memcpy(inv_evecs, p->inv_eigenvecs[0], 16 * sizeof(double));
  memcpy(evals, p->eigenvals[0], 4 * sizeof(double));
}

void ref_get_clv(void* hv, unsigned clv_index, double* out)
{
  pll_partition_t* p = static_cast<RefHandle*>(hv)->part->partition_.get();
  memcpy(out, p->clv[clv_index], sizeof(double) * p->sites * p->rate_cats * 4);
}

void ref_get_scaler(void* hv, unsigned scaler_index, unsigned* out)
{
  pll_partition_t* p = static_cast<RefHandle*>(hv)->part->partition_.get();
  memcpy(out, p->scale_buffer[scaler_index], sizeof(unsigned) * p->sites);
}

void ref_get_pmatrix(void* hv, unsigned idx, double* out)
{
  pll_partition_t* p = static_cast<RefHandle*>(hv)->part->partition_.get();
  memcpy(out, p->pmatrix[idx], sizeof(double) * p->rate_cats * 16);
}

/* runs Partition::TraversalUpdate and reports the op list it executed */
int ref_traversal_update(void* hv, unsigned node, int full, unsigned* ops_out /* 8 per op */)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  try {
    unsigned n = h->part->TraversalUpdate(h->by_index.at(node),
                                          full ? TraversalType::FULL : TraversalType::PARTIAL);
    if (ops_out)
      memcpy(ops_out, h->part->operations_.data(), n * sizeof(pll_operation_t));
    return (int)n;
  } catch (const std::exception& e) { g_error = e.what(); return -1; }
}

double ref_loglikelihood(void* hv, unsigned node, double* persite)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  return h->part->LogLikelihood(h->by_index.at(node), persite);
}

void ref_update_branch_length(void* hv, unsigned node, double len)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  h->part->UpdateBranchLength(h->by_index.at(node), len);
}

double ref_optimize_branch(void* hv, unsigned node)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  try { return h->part->OptimizeBranch(h->by_index.at(node)); }
  catch (const std::exception& e) { g_error = e.what(); return -1.0; }
}

void ref_optimize_all_branches_once(void* hv, unsigned node)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  h->part->OptimizeAllBranchesOnce(h->by_index.at(node));
}

void ref_optimize_all_branches(void* hv, unsigned node)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  h->part->OptimizeAllBranches(h->by_index.at(node));
}

void ref_optimize_branch_neighborhood(void* hv, unsigned node, int radius)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  h->part->OptimizeBranchNeighborhood(h->by_index.at(node), radius);
}

/* sumtable of the edge at `node` (CLVs must face each other) */
void ref_sumtable(void* hv, unsigned node, double* out)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  pll_partition_t* p = h->part->partition_.get();
  pll_unode_t* n = h->by_index.at(node);
  pll_update_sumtable(p, n->clv_index, n->back->clv_index, n->scaler_index,
                      n->back->scaler_index, h->part->params_indices_.data(),
                      h->part->sumtable_);
  if (out) memcpy(out, h->part->sumtable_, sizeof(double) * p->sites * p->rate_cats * 4);
}

/* derivatives of -lnL from the sumtable last computed by ref_sumtable */
void ref_derivatives(void* hv, unsigned node, double len, double* d1, double* d2)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  pll_partition_t* p = h->part->partition_.get();
  pll_unode_t* n = h->by_index.at(node);
  pll_compute_likelihood_derivatives(p, n->scaler_index, n->back->scaler_index, len,
                                     h->part->params_indices_.data(),
                                     h->part->sumtable_, d1, d2);
}

/* tree edits used by the move testers (LIBPLL/utree_moves.c:72-109,
   PTPLL/pll_util.cpp:353-366) */
int ref_nni(void* hv, unsigned node, int type)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  return pll_utree_nni(h->by_index.at(node), type, nullptr);
}

void ref_invalidate_edge(void* hv, unsigned node)
{
  RefHandle* h = static_cast<RefHandle*>(hv);
  pt::pll::InvalidateEdgeClvs(h->by_index.at(node));
}

double ref_length(void* hv, unsigned node)
{
  return static_cast<RefHandle*>(hv)->by_index.at(node)->length;
}

} // extern "C"
