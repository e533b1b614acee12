"""ptgpu_optimize_all_branches (Partition::OptimizeAllBranches as one persistent kernel)
against the same procedure issued call by call through the single-purpose entry points --
bit for bit: lengths, lnL, pass count, and every CLV / scaler / P matrix the engine is left
with -- and against the unmodified reference's result (golden `oab_lnl`, `oab_lengths`)."""
import numpy as np
import pytest

import all_branches as ab
from common import fh, load_set, make_engine, model_of
from pytree import Tree

pytestmark = pytest.mark.gpu


def fresh(name):
    tree, seqs, g = load_set(name)
    return tree, seqs, g, model_of(g)


def engine_state(e, tree):
    n = tree.tip_count
    clv = [e.get_clv(n + i) for i in range(n - 2)]
    scal = [e.get_scaler(i) for i in range(n - 2)]
    pm = [e.get_pmatrix(i) for i in range(2 * n - 3)]
    return clv, scal, pm


def run_fused(e, tree, flags):
    n_edges = 2 * tree.tip_count - 3
    steps, n_pro, n_body, n_bodies, edge, snaps = ab.program(tree.root, flags, n_edges)
    lengths0 = np.zeros(n_edges)
    for r in tree.by_index.values():
        lengths0[r.pmatrix_index] = r.length
    out, lnl, passes = e.optimize_all_branches(steps, n_pro, n_body, n_bodies, edge, lengths0)
    for r in tree.by_index.values():
        r.length = out[r.pmatrix_index]
    return lnl, passes


@pytest.mark.parametrize("name", ["ds1", "five", "newton"])
def test_fused_equals_call_by_call_and_reference(name):
    from pt_b200.engine import ops_array
    t1, seqs, g, model = fresh(name)
    t2 = Tree(t1.newick())
    e1, e2 = make_engine(t1, seqs, *model), make_engine(t2, seqs, *model)
    try:
        # the state the golden run starts from: FULL traversal at the root
        for e, t in ((e1, t1), (e2, t2)):
            ops, midx, mlen = t.full_traversal()
            e.update_prob_matrices(midx, mlen)
            e.update_partials(ops_array(ops))
        f1, f2 = ab.Flags(), ab.Flags()
        for f, t in ((f1, t1), (f2, t2)):
            ab.traverse(t.root, f, partial=False)
        lnl1, passes1 = ab.stepwise(e1, t1.root, f1)
        lnl2, passes2 = run_fused(e2, t2, f2)
        assert passes1 == passes2
        assert lnl1 == lnl2
        for pm, length in t1.branches().items():
            assert t2.branches()[pm] == length, pm
        s1, s2 = engine_state(e1, t1), engine_state(e2, t2)
        for a, b in zip(s1, s2):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
        # the reference (tests/golden/make_golden.py: OptimizeAllBranches from the input lengths)
        assert abs(lnl2 - fh(g["oab_lnl"])) < 1e-9 * abs(lnl2)
        for pm, v in g["oab_lengths"].items():
            assert abs(t2.branches()[int(pm)] - fh(v)) < 1e-6 * max(1.0, fh(v)), pm
    finally:
        e1.close()
        e2.close()


def test_fused_from_a_stale_orientation():
    """start from CLVs oriented at another node (what the wanderer's clones look like): the
    first pass then records different traversals from the later ones"""
    from pt_b200.engine import ops_array
    t1, seqs, g, model = fresh("ds1")
    t2 = Tree(t1.newick())
    e1, e2 = make_engine(t1, seqs, *model), make_engine(t2, seqs, *model)
    try:
        flags = []
        for e, t in ((e1, t1), (e2, t2)):
            other = t.nodes[t.tip_count + 7]
            ops, midx, mlen = t.full_traversal(other)
            e.update_prob_matrices(midx + [other.pmatrix_index], mlen + [other.length])
            e.update_partials(ops_array(ops))
            f = ab.Flags()
            ab.traverse(other, f, partial=False)
            flags.append(f)
        lnl1, passes1 = ab.stepwise(e1, t1.root, flags[0])
        lnl2, passes2 = run_fused(e2, t2, flags[1])
        assert (lnl1, passes1) == (lnl2, passes2)
        assert t1.branches() == t2.branches()
        for a, b in zip(engine_state(e1, t1), engine_state(e2, t2)):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
    finally:
        e1.close()
        e2.close()


def test_invalid_all_branches_programs_are_rejected():
    from pt_b200.engine import PtGpuError, ops_array
    tree, seqs, g, model = fresh("five")
    e = make_engine(tree, seqs, *model)
    try:
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        flags = ab.Flags()
        ab.traverse(tree.root, flags, partial=False)
        n_edges = 2 * tree.tip_count - 3
        steps, n_pro, n_body, n_bodies, edge, _ = ab.program(tree.root, flags, n_edges)
        l0 = np.zeros(n_edges)
        for r in tree.by_index.values():
            l0[r.pmatrix_index] = r.length

        def refuse(**kw):
            args = dict(steps=steps.copy(), edge=edge.copy(), l0=l0.copy())
            for k, f in kw.items():
                f(args[k])
            with pytest.raises(PtGpuError):
                e.optimize_all_branches(args["steps"], n_pro, n_body, n_bodies, args["edge"], args["l0"])

        k_op = int(np.nonzero(steps["type"] == 1)[0][0])
        k_opt = int(np.nonzero(steps["type"] == 2)[0][0])
        refuse(steps=lambda s: s["a_clv"].__setitem__(k_op, 0))              # CLV update into a tip
        refuse(steps=lambda s: s["a_clv"].__setitem__(k_op, 10 ** 6))
        refuse(steps=lambda s: s["b_matrix"].__setitem__(k_opt, n_edges))     # matrix out of range
        refuse(steps=lambda s: s["length_slot"].__setitem__(k_opt, 0 if s["b_matrix"][k_opt] else 1))
        refuse(steps=lambda s: s["type"].__setitem__(k_opt, 7))
        refuse(edge=lambda ed: ed["matrix"].__setitem__(0, n_edges))
        refuse(l0=lambda x: x.__setitem__(0, float("nan")))
        out, lnl, passes = e.optimize_all_branches(steps, n_pro, n_body, n_bodies, edge, l0)
        assert np.isfinite(lnl) and passes >= 1
    finally:
        e.close()
