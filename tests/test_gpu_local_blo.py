"""Batched radius >= 1 NNI scoring (ptgpu_optimize_local_batch) against (a) golden
vectors of the unmodified reference's BranchNeighborhoodOptimizer and (b) the same
procedure driven call by call through the single-edge C-ABI entry points, which
must agree bit for bit (same kernels' arithmetic, same reduction trees)."""
import json
import os

import numpy as np
import pytest

from common import GOLD, fh, load_set, make_engine, model_of
from pytree import Tree, random_tree_newick

pytestmark = pytest.mark.gpu


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def batch_engine(tree, seqs, model):
    from pt_b200.engine import ops_array
    e = make_engine(tree, seqs, *model, directional=True)
    br = tree.branches()
    e.update_prob_matrices(list(br.keys()), list(br.values()))
    e.update_partials(ops_array(tree.directional_plan()))
    return e


def program(tree, radius):
    from pt_b200.local_blo import Program
    n = tree.tip_count
    return Program(tree, radius, n, 3 * (n - 2), 3 * (n - 2), 2 * n - 3)


def sequential(e, tree, p, kind, radius, smoothings=32, tolerance=1e-6):
    """BranchNeighborhoodOptimizer::EvaluateMove through the single-edge entry points
    on a libpll-layout engine (one CLV buffer per inner node); returns
    (lnl, {pmatrix: length})."""
    from pt_b200.engine import ops_array
    from pt_b200.local_blo import apply_nni
    saved = {r.pmatrix_index: r.length for r in tree.by_index.values()}
    apply_nni(p, kind)
    ops, midx, mlen = tree.full_traversal(p)
    e.update_prob_matrices(midx + [p.pmatrix_index], mlen + [p.length])
    e.update_partials(ops_array(ops))
    xmin, xmax, xtol = 1e-4, 100.0, 1e-5

    def op(parent, c1, c2):
        b, c = c1.back, c2.back
        e.update_partials(ops_array([(parent.clv_index, parent.scaler_index, b.clv_index,
                                      b.pmatrix_index, b.scaler_index, c.clv_index,
                                      c.pmatrix_index, c.scaler_index)]))

    def recomp(r, radius):
        guess = r.length if xmin <= r.length <= xmax else 0.1
        ln, status, _ = e.optimize_branch_bounded(r.clv_index, r.scaler_index, r.back.clv_index,
                                                  r.back.scaler_index, guess, xmin, xmax, xtol, 30)
        if status:
            return False
        if abs(r.length - ln) > 1e-10:
            e.update_prob_matrices([r.pmatrix_index], [ln])
        r.length = r.back.length = ln
        q = r.next
        if radius and q is not None:
            z = q.next
            op(q, r, z)
            if not recomp(q.back, radius - 1):
                return False
            op(z, q, r)
            if not recomp(z.back, radius - 1):
                return False
            op(r, z, q)
        return True

    def edge(a, b):
        return e.edge_loglikelihood(a.clv_index, a.scaler_index, b.clv_index, b.scaler_index,
                                    a.pmatrix_index)

    q = p.back
    lnl = edge(q, p)
    it = smoothings
    while it:
        if not (recomp(p, radius) and recomp(q, radius - 1)):
            break
        now = edge(q, p)
        if not (now - lnl > now * 1e-14):
            lnl = now
            break
        it -= 1
        if abs(now - lnl) < tolerance:
            it = 0
        lnl = now
    score = edge(p, q)
    lengths = {r.pmatrix_index: r.length for r in tree.by_index.values()
               if r.length != saved[r.pmatrix_index]}
    for r in tree.by_index.values():
        r.length = saved[r.pmatrix_index]
    apply_nni(p, kind)
    for r in tree.by_index.values():
        r.length = saved[r.pmatrix_index]
    return score, lengths


@pytest.mark.parametrize("name", ["ds1", "five"])
@pytest.mark.parametrize("radius", [1, 2, 3])
def test_against_reference_golden(name, radius):
    with open(os.path.join(GOLD, "neighborhood_golden.json")) as f:
        gold = json.load(f)[name][str(radius)]
    tree, seqs, g = load_set(name)
    e = batch_engine(tree, seqs, model_of(g))
    try:
        prog = program(tree, radius)
        lens, lnl, status, sm = e.optimize_local_batch(prog)
        assert prog.count == len(gold)
        for i, row in enumerate(gold):
            assert prog.moves[i] == (row["node"], row["type"])
            assert status[i] == 0
            assert rel(lnl[i], fh(row["lnl"])) < 1e-10, (i, lnl[i], fh(row["lnl"]))
            edges = prog.slot_edges[i]
            got = {edges[s]: lens[i, s] for s in edges}
            for pm, v in row["lengths"].items():
                assert abs(got[int(pm)] - fh(v)) < 1e-7 * max(1.0, fh(v)), (i, pm)
    finally:
        e.close()


@pytest.mark.parametrize("name,radius", [("ds1", 1), ("ds1", 2), ("five", 3)])
def test_batch_is_bit_identical_to_single_edge_calls(name, radius):
    tree, seqs, g = load_set(name)
    model = model_of(g)
    e = batch_engine(tree, seqs, model)
    s = make_engine(tree, seqs, *model)
    try:
        prog = program(tree, radius)
        lens, lnl, status, sm = e.optimize_local_batch(prog)
        cands = [(p, k) for p in tree.nodes[tree.tip_count:] if not p.back.is_tip for k in (1, 2)]
        assert len(cands) == prog.count
        for i, (p, kind) in enumerate(cands):
            score, changed = sequential(s, tree, p, kind, radius)
            assert score == lnl[i], (i, score, lnl[i])
            edges = prog.slot_edges[i]
            got = {edges[k]: lens[i, k] for k in edges}
            for pm, v in changed.items():
                assert got[pm] == v, (i, pm)
    finally:
        e.close()
        s.close()


@pytest.mark.parametrize("radius", [0, 1, 2])
def test_pruned_directional_plan_gives_identical_scores(radius):
    """only the directional CLVs a batch reads need to be current: never one that faces a tip"""
    from pt_b200.engine import ops_array
    from pytree import cand_array
    tree, seqs, g = load_set("ds1")
    model = model_of(g)
    results = []
    for pruned in (False, True):
        e = make_engine(tree, seqs, *model, directional=True)
        try:
            br = tree.branches()
            e.update_prob_matrices(list(br.keys()), list(br.values()))
            if radius == 0:
                cands = tree.nni_candidates()
                plan = tree.directional_plan(tree.referenced_clvs(cands) if pruned else None)
                e.update_partials(ops_array(plan))
                lens, lnl, status = e.score_nni_batch(cand_array(cands))
            else:
                prog = program(tree, radius)
                plan = tree.directional_plan(prog.referenced_clvs if pruned else None)
                e.update_partials(ops_array(plan))
                lens, lnl, status, _ = e.optimize_local_batch(prog)
            results.append((len(plan), lens.copy(), lnl.copy(), status.copy()))
        finally:
            e.close()
    full, pruned = results
    assert pruned[0] < full[0] == 3 * (tree.tip_count - 2)
    assert all(np.array_equal(a, b) for a, b in zip(full[1:], pruned[1:]))


@pytest.mark.parametrize("knob", ["PTGPU_BLO_MAX_CHUNK=7", "PTGPU_BLO_CLUSTER_WORK=0",
                                  "PTGPU_BLO_CLUSTER_WORK=100000"])
def test_launch_geometry_does_not_change_results(knob, monkeypatch):
    """the same batch in chunks of 7 candidates (private indices rebased per chunk), forced onto
    the cooperative grid-barrier variant, and forced onto the single-cluster variant: a
    candidate's numbers depend on neither its batch nor the barrier mechanism"""
    tree, seqs, g = load_set("ds1")
    model = model_of(g)
    prog = program(tree, 2)
    e = batch_engine(tree, seqs, model)
    try:
        base = e.optimize_local_batch(prog)
    finally:
        e.close()
    name, value = knob.split("=")
    monkeypatch.setenv(name, value)
    e = batch_engine(tree, seqs, model)
    try:
        got = e.optimize_local_batch(prog)
    finally:
        e.close()
    for a, b in zip(base, got):
        assert np.array_equal(a, b)


def test_batch_on_a_larger_alignment_matches_sampled_sequential_runs():
    import synth
    taxa, sites = 24, 20000
    tree = synth.parent_tree(taxa, 0)
    seqs = {k: bytes(v).decode() for k, v in
            synth.simulate(tree, sites, np.random.default_rng(synth.SEED_ALN)).items()}
    model = (synth.FREQS, synth.SUBST, synth.RATES)
    e = batch_engine(tree, seqs, model)
    s = make_engine(tree, seqs, *model)
    try:
        prog = program(tree, 1)
        lens, lnl, status, sm = e.optimize_local_batch(prog)
        assert (status == 0).all() and np.isfinite(lnl).all()
        cands = [(p, k) for p in tree.nodes[tree.tip_count:] if not p.back.is_tip for k in (1, 2)]
        for i in (0, 7, len(cands) - 1):
            score, _ = sequential(s, tree, *cands[i], 1)
            assert score == lnl[i]
        # a candidate's numbers do not depend on its batch: re-run a sub-batch
        again = e.optimize_local_batch(prog)
        assert np.array_equal(again[1], lnl) and np.array_equal(again[0], lens)
    finally:
        e.close()
        s.close()


def test_invalid_programs_are_rejected():
    """a local batch may write only its own private slots; every index is checked on the host"""
    from pt_b200.engine import PtGpuError
    tree, seqs, g = load_set("five")
    e = batch_engine(tree, seqs, model_of(g))
    n = tree.tip_count
    try:
        good = program(tree, 1)
        e.optimize_local_batch(good)

        def broken(edit):
            prog = program(tree, 1)
            edit(prog)
            with pytest.raises(PtGpuError):
                e.optimize_local_batch(prog)

        def write_engine_clv(p):       # CLV update into one of the engine's directional buffers
            p.steps["a_clv"][0, 0] = n
        def write_other_candidate(p):  # ... into another candidate's private buffer
            p.steps["a_clv"][0, 0] = p.steps["a_clv"][0, 1]
        def foreign_matrix(p):         # optimise an edge whose matrix is not the slot's own
            k = int(np.nonzero(p.step_types == 2)[0][0])
            p.steps["b_matrix"][k, 0] = 0
        def slot_out_of_range(p):
            k = int(np.nonzero(p.step_types == 2)[0][0])
            p.steps["length_slot"][k, 0] = p.n_lengths
        def wrong_type(p):
            p.steps["type"][0, 0] = 2
        def negative_length(p):
            p.lengths0[0, 0] = -1.0
        def bad_edge(p):
            p.final_edges["child_clv"][0] = 10 ** 6
        for edit in (write_engine_clv, write_other_candidate, foreign_matrix, slot_out_of_range,
                     wrong_type, negative_length, bad_edge):
            broken(edit)
        # the engine still works after the refusals
        again = e.optimize_local_batch(good)
        assert np.isfinite(again[1]).all()
    finally:
        e.close()
