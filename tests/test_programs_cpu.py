"""Host-side (integer) logic of the step programs, no GPU needed: the unrolled
recomp_iterative recursion (pt_b200/local_blo.py) and the recorded OptimizeAllBranches
procedure (tests/all_branches.py) are checked for shape, index ranges and the invariants
the kernels rely on."""
import numpy as np
import pytest

import all_branches as ab
from common import load_set
from pt_b200.local_blo import (BLO_MAT, BLO_NOP, BLO_OP, BLO_OPT, Program, apply_nni, template,
                               template_recomp)
from pytree import Tree, random_tree_newick


def ds1():
    return load_set("ds1")[0]


@pytest.mark.parametrize("radius", [1, 2, 3, 4])
def test_template_sizes_follow_the_recursion(radius):
    # recomp(r) = OPT + [OP recomp(r-1) OP recomp(r-1) OP]; a smoothing = recomp(r) + recomp(r-1)
    size = lambda r: 1 if r == 0 else 4 + 2 * size(r - 1)
    assert len(template_recomp(radius)) == size(radius)
    assert len(template(radius)) == size(radius) + size(radius - 1)
    assert template(radius).count(BLO_OPT) == sum(1 for t in template(radius) if t == BLO_OPT)


def test_nni_is_its_own_inverse():
    t = ds1()
    before = t.newick()
    for p in t.nodes[t.tip_count:]:
        if p.back.is_tip:
            continue
        for kind in (1, 2):
            apply_nni(p, kind)
            assert t.newick() != before
            apply_nni(p, kind)
            assert t.newick() == before


@pytest.mark.parametrize("radius", [1, 2, 3])
@pytest.mark.parametrize("seed", [1, 2])
def test_program_invariants(radius, seed):
    rng = np.random.default_rng(seed)
    t = Tree(random_tree_newick(20, rng))
    n = t.tip_count
    nbuf, nmat = 3 * (n - 2), 2 * n - 3
    newick = t.newick()
    P = Program(t, radius, n, nbuf, nbuf, nmat)
    assert t.newick() == newick                       # the tree is handed back untouched
    assert P.count == len(t.nni_candidates())
    assert P.steps.shape == (2 + 2 * P.n_body, P.count)
    assert list(P.step_types[:2]) == [BLO_OP, BLO_OP]
    assert list(P.step_types[2:2 + P.n_body]) == list(P.step_types[2 + P.n_body:]) == template(radius)
    clv_total = n + nbuf
    for i in range(P.count):
        col = P.steps[:, i]
        lo, hi = clv_total + i * P.n_private_clvs, clv_total + (i + 1) * P.n_private_clvs
        mlo, mhi = nmat + i * P.n_lengths, nmat + (i + 1) * P.n_lengths
        written = set()
        for k, st in enumerate(col):
            assert st["type"] in (BLO_NOP, P.step_types[k])
            if st["type"] == BLO_OP:
                assert lo <= st["a_clv"] < hi                      # writes stay private
                assert st["a_scaler"] == nbuf + (st["a_clv"] - clv_total)
                written.add(int(st["a_clv"]))
                for side in ("b", "c"):
                    c = int(st[side + "_clv"])
                    assert c < clv_total or c in written            # private CLVs are read after being written
                    assert st[side + "_matrix"] < nmat or mlo <= st[side + "_matrix"] < mhi
            elif st["type"] == BLO_OPT:
                assert st["b_matrix"] == nmat + i * P.n_lengths + st["length_slot"]
                for side in ("a", "b"):
                    c = int(st[side + "_clv"])
                    assert c < clv_total or c in written
        # the second body variant never reads an engine CLV of a node it has a private copy of
        private_nodes = {int(s["a_clv"]) for s in col if s["type"] == BLO_OP}
        assert len(private_nodes) <= P.n_private_clvs
        for e in (P.smooth_edges[i], P.final_edges[i]):
            assert lo <= e["parent_clv"] < hi and lo <= e["child_clv"] < hi
            assert e["matrix"] == nmat + i * P.n_lengths           # slot 0 = the central branch
        assert P.smooth_edges[i]["parent_clv"] == P.final_edges[i]["child_clv"]
    # lengths: slot 0 is the central branch of the candidate
    for i, (node_index, kind) in enumerate(P.moves):
        assert P.lengths0[i, 0] == t.by_index[node_index].length


def test_referenced_directional_clvs_never_face_a_tip():
    t = ds1()
    n = t.tip_count
    for refs in (t.referenced_clvs(t.nni_candidates()),
                 Program(t, 2, n, 3 * (n - 2), 3 * (n - 2), 2 * n - 3).referenced_clvs):
        for idx in refs:
            if idx >= n:
                assert not t.by_index[idx].back.is_tip
        plan = t.directional_plan(refs)
        produced = {row[0] for row in plan}
        assert {i for i in refs if i >= n} <= produced
        assert len(plan) < 3 * (n - 2)
        # dependency order: children are produced before they are consumed
        seen = set()
        for row in plan:
            for child in (row[2], row[5]):
                assert child < n or child in seen
            seen.add(row[0])


def test_all_branches_program_reaches_a_fixed_point_and_matches_the_call_sequence():
    t = ds1()
    n_edges = 2 * t.tip_count - 3
    flags = ab.Flags()
    ab.traverse(t.root, flags, partial=False)
    steps, n_pro, n_body, n_bodies, edge, snaps = ab.program(t.root, flags, n_edges)
    assert n_pro == 0 and n_bodies == 2            # CLVs already face the root: nothing to do first
    body0 = steps[n_pro:n_pro + n_body]
    body1 = steps[n_pro + n_body:n_pro + 2 * n_body]
    nodes = 2 * t.tip_count - 2
    assert (body0["type"] == BLO_OPT).sum() == (body1["type"] == BLO_OPT).sum() == nodes
    assert (body0["type"] == BLO_MAT).sum() == n_edges        # every matrix refreshed once ...
    assert (body1["type"] == BLO_MAT).sum() == 0              # ... and never again
    assert snaps[0] == snaps[1]
    # every optimisation is preceded by the update of the node it is rooted at
    last_written = None
    for st in body0:
        if st["type"] == BLO_OP:
            last_written = int(st["a_clv"])
        elif st["type"] == BLO_OPT and last_written is not None:
            assert st["a_clv"] == last_written or st["a_clv"] >= t.tip_count
    assert edge[0]["parent_clv"] == t.root.clv_index
