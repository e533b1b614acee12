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
