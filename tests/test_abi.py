"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every
symbol include/ptgpu.h declares; without a CUDA device calls fail loudly (there
is no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ptgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ptgpu_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_the_hot_path():
    syms = declared_symbols()
    for s in ("ptgpu_create", "ptgpu_update_prob_matrices", "ptgpu_update_partials",
              "ptgpu_compute_edge_loglikelihood", "ptgpu_update_sumtable",
              "ptgpu_compute_likelihood_derivatives", "ptgpu_optimize_branch",
              "ptgpu_score_nni_batch"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    from pt_b200.engine import LIB_PATH, load_library
    assert os.path.exists(LIB_PATH), "run __graft_entry__.build() first"
    lib = load_library()
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} missing from libptgpu.so"


def test_struct_layouts_match_libpll():
    from pt_b200.engine import CAND_DTYPE, OP_DTYPE
    # pll_operation_t: 8 x 4-byte fields in this order (libpll pll.h:249-259)
    assert OP_DTYPE.names == ("parent_clv_index", "parent_scaler_index", "child1_clv_index",
                              "child1_matrix_index", "child1_scaler_index", "child2_clv_index",
                              "child2_matrix_index", "child2_scaler_index")
    assert OP_DTYPE.itemsize == 32 and CAND_DTYPE.itemsize == 72


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from pt_b200.engine import Engine, PtGpuError
    with pytest.raises(PtGpuError):
        Engine(4, 2, 10, 4, 5, 2)


def test_product_does_not_touch_the_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pt_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dirpath, fn), errors="ignore").read()
                if re.search(r"pyoracle|pt_oracle|libpllref|oracle/", txt):
                    bad.append(fn)
    assert not bad, bad


def test_gamma_rates_host_routine_matches_reference_golden():
    """ptgpu_gamma_rates_mean replaces pll_compute_gamma_cats (host-only, no device needed):
    bit-identical to the rates the reference computed for DS1 / five / newton"""
    import numpy as np
    from common import fha, golden
    from pt_b200.engine import load_library
    lib = load_library()
    lib.ptgpu_gamma_rates_mean.argtypes = [ctypes.c_double, ctypes.c_uint,
                                           ctypes.POINTER(ctypes.c_double)]
    for name, alpha in (("ds1", 0.134946),):
        g = golden(name)
        out = np.empty(4)
        rc = lib.ptgpu_gamma_rates_mean(alpha, 4, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
        assert rc == 0
        assert np.array_equal(out, fha(g["rates"]))
    out = np.empty(1)
    assert lib.ptgpu_gamma_rates_mean(0.5, 1, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double))) == 0
    assert out[0] == 1.0
    assert lib.ptgpu_gamma_rates_mean(0.01, 4, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double))) != 0


def test_gamma_rates_against_oracle_grid(oracle):
    import numpy as np
    from pt_b200.engine import load_library
    lib = load_library()
    lib.ptgpu_gamma_rates_mean.argtypes = [ctypes.c_double, ctypes.c_uint,
                                           ctypes.POINTER(ctypes.c_double)]
    for alpha in (0.02, 0.05, 0.1299395, 0.3, 0.5, 1.0, 2.5, 10.0, 99.0):
        for cats in (2, 4, 8):
            out = np.empty(cats)
            assert lib.ptgpu_gamma_rates_mean(alpha, cats, out.ctypes.data_as(
                ctypes.POINTER(ctypes.c_double))) == 0
            assert np.array_equal(out, oracle.gamma_rates_mean(alpha, cats)), (alpha, cats)


def test_host_layer_mirrors_the_reference_surface():
    """every public method of the reference's pt::pll::Partition is declared by the host layer"""
    hdr = open(os.path.join(ROOT, "pt_b200", "host", "include", "pll_partition.hpp")).read()
    for m in ("LogLikelihood(pll_unode_t* tree, double* per_site_lnl = nullptr)",
              "LogMarginalLikelihood(pll_unode_t* root)",
              "TraversalUpdate(pll_unode_t* root, TraversalType type)",
              "UpdateBranchLength(pll_unode_t* node, double length)",
              "OptimizeBranch(pll_unode_t* node)",
              "OptimizeAllBranchesOnce(pll_unode_t* tree, bool use_pllmod = false)",
              "OptimizeAllBranches(pll_unode_t* tree, bool use_pllmod = false)",
              "OptimizeBranchNeighborhood(pll_unode_t* node, int radius)",
              "OptimizeModelOnce(pll_unode_t* tree)", "OptimizeModel(pll_unode_t* tree)",
              "OptimizeAllBranchesAndModel(pll_unode_t* tree, bool use_pllmod = false)",
              "Model GetModel() const", "void SetModel(const Model& model)",
              "tip_node_count() const", "inner_node_count() const", "node_count() const",
              "branch_count() const", "const unsigned int MAX_ITER = 32", "const double EPSILON = 1e-6"):
        assert m in hdr, m
