"""ptgpu_compute_edge_loglikelihood_noalloc: the edge log-likelihood with its scratch in a
file-scope device array of the library (no allocation of any kind in the call; no counterpart
in the reference).  Through the C ABI, against the existing
entry point, on two calls in a row; and the size limit.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from common import fh, load_set, make_engine, model_of

pytestmark = pytest.mark.gpu


def noalloc_lnl(e, pclv, pscaler, cclv, cscaler, matrix):
    out = C.c_double()
    rc = e.lib.ptgpu_compute_edge_loglikelihood_noalloc(
        e.h, C.c_uint(pclv), C.c_int(pscaler), C.c_uint(cclv), C.c_int(cscaler), C.c_uint(matrix),
        C.byref(out))
    return rc, out.value


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_noalloc_edge_lnl_equals_the_allocating_entry_point_twice_in_a_row(name):
    from pt_b200.engine import ops_array
    tree, seqs, g = load_set(name)
    e = make_engine(tree, seqs, *model_of(g))
    try:
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        root = tree.by_index[g["root"]]
        rb = root.back
        edge = (root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index, root.pmatrix_index)
        want = e.edge_loglikelihood(*edge)
        assert abs(want - fh(g["lnl"])) <= 1e-12 * abs(want)
        rc, got = noalloc_lnl(e, *edge)
        assert rc == 0, e.lib.ptgpu_last_error()
        assert abs(got - want) <= 1e-12 * abs(want) and got == want    # same kernel, same reduction tree
        # second call in a row on another state: the shared scratch holds the first call's leftovers
        e.update_prob_matrices([root.pmatrix_index], [root.length * 1.7 + 0.01])
        want2 = e.edge_loglikelihood(*edge)
        rc, got2 = noalloc_lnl(e, *edge)
        assert rc == 0, e.lib.ptgpu_last_error()
        assert want2 != want and abs(got2 - want2) <= 1e-12 * abs(want2) and got2 == want2
        # and back: a third call reproduces the first
        e.update_prob_matrices([root.pmatrix_index], [root.length])
        rc, got3 = noalloc_lnl(e, *edge)
        assert rc == 0 and got3 == want
        # bad indices are rejected like the other entry point's
        rc, _ = noalloc_lnl(e, 10 ** 6, -1, rb.clv_index, rb.scaler_index, root.pmatrix_index)
        assert rc != 0
    finally:
        e.close()


def test_noalloc_rejects_engines_beyond_the_fixed_scratch():
    from pt_b200.engine import Engine, load_library
    lib = load_library()
    lib.ptgpu_noalloc_max_sites.restype = C.c_uint
    limit = int(lib.ptgpu_noalloc_max_sites())
    assert limit == 8192 * 256
    e = Engine(3, 1, limit + 1, 4, 3, 1)
    try:
        e.set_model([0.25] * 4, [1.0] * 6, [1.0] * 4)
        rc, _ = noalloc_lnl(e, 3, 0, 0, -1, 0)
        assert rc == -2 and b"exceed the fixed scratch" in lib.ptgpu_last_error()
    finally:
        e.close()
