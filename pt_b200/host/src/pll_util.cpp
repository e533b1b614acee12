// pt_b200/host/src/pll_util.cpp -- host-side tree bookkeeping (no likelihood
// arithmetic).  Behaviour follows the reference's lib/libptpll/src/pll_util.cpp
// (cited per function as PTPLL/pll_util.cpp:line); the code is written for this
// repository.
#include "pll_util.hpp"

#include <cstdlib>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace pt { namespace pll {

namespace {

node_info_t* info(pll_unode_t* n) { return static_cast<node_info_t*>(n->data); }

node_info_t* new_info(int valid)
{
  node_info_t* p = static_cast<node_info_t*>(std::malloc(sizeof(node_info_t)));
  if (!p) throw std::bad_alloc();
  p->clv_valid = valid;
  return p;
}

// Orientation protocol of the shared per-node CLV buffer
// (PTPLL/pll_util.cpp:45-94): an inner node has ONE CLV buffer; the record
// whose flag is set is the direction the buffer currently serves.  Returning
// 1 puts the node into the traversal (its CLV gets recomputed), 0 prunes the
// subtree.
int orient(pll_unode_t* node, TraversalType type)
{
  if (!node->next) return 1;                      // tips always enter the list
  pll_unode_t* a = node->next;
  pll_unode_t* b = a->next;
  if (!node->data) {
    node->data = new_info(1);
    a->data = new_info(0);
    b->data = new_info(0);
    return 1;
  }
  if (type == TraversalType::PARTIAL && info(node)->clv_valid) return 0;
  info(node)->clv_valid = 1;
  info(a)->clv_valid = 0;
  info(b)->clv_valid = 0;
  return 1;
}

} // namespace

// every record of a node must carry a non-zero length equal on both ends of
// its edge (PTPLL/pll_util.cpp:25-34)
int cb_branch_healthy(pll_unode_t* tree)
{
  if (!tree->length) return 0;
  if (tree->next && (!tree->next->length || !tree->next->next->length)) return 0;
  return tree->length == tree->back->length;
}

bool TreeHealthy(pll_utree_t* tree)
{
  return pll_utree_every(tree, cb_branch_healthy);
}

int cb_full_traversal(pll_unode_t* node) { return orient(node, TraversalType::FULL); }
int cb_partial_traversal(pll_unode_t* node) { return orient(node, TraversalType::PARTIAL); }

// after pll_utree_clone the three records of a node still point at the
// SOURCE tree's flags; give the clone its own copies (PTPLL/pll_util.cpp:109-139)
int cb_copy_clv_traversal(pll_unode_t* node)
{
  if (!node->next || !node->data) return 1;
  pll_unode_t* ring[3] = {node, node->next, node->next->next};
  int valid[3];
  for (int k = 0; k < 3; ++k) valid[k] = info(ring[k])->clv_valid;
  for (int k = 0; k < 3; ++k) ring[k]->data = new_info(valid[k]);
  return 1;
}

void cb_erase_data(void* data) { std::free(data); }

// FASTA reader (PTPLL/pll_util.cpp:240-302): `seq_count` records of equal
// length; returns the number of sites.
unsigned int ParseFasta(std::string path, unsigned int seq_count,
                        std::vector<std::string>& headers_out,
                        std::vector<std::string>& seqdata_out)
{
  pll_fasta_t* fp = pll_fasta_open(path.c_str(), pll_map_fasta);
  if (!fp) throw std::invalid_argument("Error opening file " + path);

  headers_out.assign(seq_count, std::string());
  seqdata_out.assign(seq_count, std::string());

  char* hdr = nullptr;
  char* seq = nullptr;
  long hdrlen = 0, seqlen = 0, seqno = 0;
  long sites = -1;
  unsigned int n = 0;
  while (pll_fasta_getnext(fp, &hdr, &hdrlen, &seq, &seqlen, &seqno)) {
    std::string h(hdr), s(seq);
    std::free(hdr);
    std::free(seq);
    if (n >= seq_count) {
      pll_fasta_close(fp);
      throw std::invalid_argument("FASTA file contains more sequences than expected");
    }
    if (sites >= 0 && sites != seqlen) {
      pll_fasta_close(fp);
      throw std::invalid_argument("FASTA file does not contain equal size sequences");
    }
    sites = seqlen;
    headers_out[n].swap(h);
    seqdata_out[n].swap(s);
    ++n;
  }
  pll_fasta_close(fp);

  if (pll_errno != PLL_ERROR_FILE_EOF)
    throw std::runtime_error("Error while reading file " + path);
  if (sites < 0) throw std::runtime_error("Unable to read alignment");
  if (n != seq_count) throw std::invalid_argument("Some taxa are missing from FASTA file");
  return static_cast<unsigned int>(sites);
}

// one Newick tree per non-empty line; tip indices of trees 1.. are rewritten
// to tree 0's so one partition serves all of them (PTPLL/pll_util.cpp:304-346)
std::vector<pll_utree_t*> ParseMultiNewick(const std::string& filename)
{
  std::ifstream in(filename);
  if (!in) throw std::invalid_argument("Error opening file " + filename);

  std::vector<pll_utree_t*> trees;
  std::string line;
  while (std::getline(in, line)) {
    if (line.empty()) continue;
    pll_utree_t* t = pll_utree_parse_newick_string(line.c_str());
    if (!t) {
      for (pll_utree_t* x : trees) pll_utree_destroy(x, nullptr);
      throw std::runtime_error("Failed to parse '" + line + "' as a Newick string");
    }
    trees.push_back(t);
  }
  for (size_t i = 1; i < trees.size(); ++i) SynchronizeTipIndices(trees[0], trees[i]);
  return trees;
}

pll_unode_t* GetVirtualRoot(pll_utree_t* tree)
{
  return tree->nodes[tree->tip_count + tree->inner_count - 1];
}

// an NNI across `node`'s edge changes what both end nodes see
// (PTPLL/pll_util.cpp:353-366)
void InvalidateEdgeClvs(pll_unode_t* node)
{
  if (node->data) info(node)->clv_valid = 0;
  if (node->back->data) info(node->back)->clv_valid = 0;
  // (this thread's latest invalidations: Partition::DeferTesterOrientation replays them)
  std::vector<pll_unode_t*>& log = InvalidationLog();
  if (log.size() >= 64) log.erase(log.begin(), log.begin() + 32);
  log.push_back(node);
}

std::vector<pll_unode_t*>& InvalidationLog()
{
  thread_local std::vector<pll_unode_t*> log;
  return log;
}

// PTPLL/pll_util.cpp:368-411
void SynchronizeTipIndices(pll_utree_t* src_tree, pll_utree_t* dest_tree)
{
  if (src_tree->tip_count != dest_tree->tip_count)
    throw std::invalid_argument("Number of tips differs in source and destination");

  std::map<std::string, pll_unode_t*> by_label;
  for (unsigned int i = 0; i < src_tree->tip_count; ++i) {
    pll_unode_t* tip = src_tree->nodes[i];
    if (!by_label.emplace(tip->label, tip).second)
      throw std::invalid_argument(std::string("Error inserting tip label ") + tip->label +
                                  " into map (possibly a duplicate)");
  }
  for (unsigned int i = 0; i < dest_tree->tip_count; ++i) {
    pll_unode_t* dst = dest_tree->nodes[i];
    auto it = by_label.find(dst->label);
    if (it == by_label.end())
      throw std::invalid_argument(std::string("Tip with label ") + dst->label +
                                  " does not appear in source tree");
    const pll_unode_t* src = it->second;
    dst->node_index = src->node_index;
    dst->clv_index = src->clv_index;
    dst->scaler_index = src->scaler_index;
    dst->pmatrix_index = src->pmatrix_index;
    dst->back->pmatrix_index = src->back->pmatrix_index;
  }
  if (!pll_utree_check_integrity(dest_tree))
    throw std::logic_error("Destination tree failed integrity check");
}

} } // namespace pt::pll
