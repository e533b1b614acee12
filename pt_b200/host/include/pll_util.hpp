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

int cb_full_traversal(pll_unode_t* node);
int cb_partial_traversal(pll_unode_t* node);
int cb_copy_clv_traversal(pll_unode_t* node);
int cb_branch_healthy(pll_unode_t* tree);
void cb_erase_data(void* data);

bool TreeHealthy(pll_utree_t* tree);

unsigned int ParseFasta(std::string path, unsigned int seq_count,
                        std::vector<std::string>& headers_out,
                        std::vector<std::string>& seqdata_out);

std::vector<pll_utree_t*> ParseMultiNewick(const std::string& filename);

pll_unode_t* GetVirtualRoot(pll_utree_t* tree);

void InvalidateEdgeClvs(pll_unode_t* node);

void SynchronizeTipIndices(pll_utree_t* src_tree, pll_utree_t* dest_tree);

} } // namespace pt::pll

#endif // PT_B200_HOST_PLL_UTIL_HPP_
