// pt_b200/host/include/pll_util.hpp -- host-side tree bookkeeping around the
// GPU engine.  Same names and meaning as the reference's
// lib/libptpll/src/pll_util.hpp:16-44, minus pllext_partition_clone (it copies
// pll_partition_t fields, has no caller in pt, and the engine's state lives in
// HBM; SURVEY.md section 8b "legacy surface with no caller").
//
// The tree itself stays libpll's pointer-linked pll_utree_t / pll_unode_t
// (host-only code that pt keeps unchanged); nothing here does likelihood
// arithmetic.
#ifndef PT_B200_HOST_PLL_UTIL_HPP_
#define PT_B200_HOST_PLL_UTIL_HPP_

#include <string>
#include <vector>

#include <libpll/pll.h>

namespace pt { namespace pll {

const unsigned int STATES = 4;

// per-record flag: is the CLV buffer of this inner node currently oriented
// towards this record's `back`? (malloc'd lazily, one per record)
typedef struct { int clv_valid; } node_info_t;

enum class TraversalType { FULL, PARTIAL };

// ---- callbacks for pll_utree_traverse / pll_utree_every ------------------------
// The two traversal callbacks implement the orientation protocol of the shared per-node
// CLV buffer (PTPLL/pll_util.cpp:45-105): return 1 = the node's CLV is recomputed and the
// flag moves to this record, 0 = the subtree behind it is pruned from the traversal.
int cb_full_traversal(pll_unode_t* node);      // every inner node is recomputed
int cb_partial_traversal(pll_unode_t* node);   // only nodes whose flag is on another record
// after pll_utree_clone the records still share the source tree's flags: give each
// record of the clone a copy of its own (PTPLL/pll_util.cpp:109-139)
int cb_copy_clv_traversal(pll_unode_t* node);
// non-zero lengths that agree on both ends of every edge of the node (:25-34)
int cb_branch_healthy(pll_unode_t* tree);
// destructor callback for pll_utree_destroy: frees one node_info_t
void cb_erase_data(void* data);

// cb_branch_healthy over the whole tree
bool TreeHealthy(pll_utree_t* tree);

// ---- input ---------------------------------------------------------------------
// `seq_count` FASTA records of equal length; returns the number of sites and throws
// std::invalid_argument on a missing file, too many / too few records or ragged lengths
// (PTPLL/pll_util.cpp:240-302)
unsigned int ParseFasta(std::string path, unsigned int seq_count,
                        std::vector<std::string>& headers_out,
                        std::vector<std::string>& seqdata_out);

// one Newick tree per non-empty line; trees 1.. get tree 0's tip indices (:304-346)
std::vector<pll_utree_t*> ParseMultiNewick(const std::string& filename);

// ---- tree helpers --------------------------------------------------------------
// the last inner node of the node array: what pt uses as the traversal root (:348-351)
pll_unode_t* GetVirtualRoot(pll_utree_t* tree);

// an NNI across node's edge changes what both of its end nodes summarise (:353-366)
void InvalidateEdgeClvs(pll_unode_t* node);
// the nodes this thread has passed to InvalidateEdgeClvs lately, oldest first (at most 64 are kept)
std::vector<pll_unode_t*>& InvalidationLog();

// copies node / clv / scaler / pmatrix indices of equally labelled tips from src to dest so
// that one engine (one set of tip buffers) serves both trees (:368-411)
void SynchronizeTipIndices(pll_utree_t* src_tree, pll_utree_t* dest_tree);

} } // namespace pt::pll

#endif // PT_B200_HOST_PLL_UTIL_HPP_
