"""Pattern weights on the device (ptgpu_set_pattern_weights +
ptgpu_compress_site_patterns) against golden vectors of the UNMODIFIED reference run on
its own pll_compress_site_patterns output (tests/golden/make_patterns_golden.py), and
against the uncompressed results: same lnL / lengths within the tolerance north_star
gives (1e-9 relative on lnL), scaler count of a pattern = scaler count of each of its
sites.  Sums over patterns use our fixed reduction tree, so the bar against the
reference's sequential sum is 1e-12 relative as for the unweighted kernels."""
import json
import os

import numpy as np
import pytest

import all_branches as ab
from common import GOLD, fh, load_set, make_engine, model_of
from pytree import Tree, cand_array

pytestmark = pytest.mark.gpu


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def gold_of(name):
    with open(os.path.join(GOLD, "patterns_golden.json")) as f:
        return json.load(f)[name]


def compressed_engine(name, directional=False):
    """engine on the unique columns the ABI's own compression yields, weights set"""
    from pt_b200.engine import compress_site_patterns
    tree, seqs, g = load_set(name)
    labels = list(seqs.keys())
    cs, w = compress_site_patterns([seqs[k] for k in labels])
    cseqs = dict(zip(labels, cs))
    e = make_engine(tree, cseqs, *model_of(g), directional=directional)
    e.set_pattern_weights(w)
    return tree, seqs, cseqs, w, g, e


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_lnl_persite_and_derivatives_match_the_reference(name):
    from pt_b200.engine import ops_array
    pg = gold_of(name)
    tree, seqs, cseqs, w, g, e = compressed_engine(name)
    try:
        assert [int(x) for x in w] == pg["weights"] and cseqs == pg["sequences"]
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        root, rb = tree.root, tree.root.back
        lnl, persite = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                            rb.scaler_index, root.pmatrix_index, persite=True)
        assert rel(lnl, fh(pg["lnl"])) < 1e-12
        assert rel(lnl, fh(pg["lnl_uncompressed"])) < 1e-9      # north_star's bar
        ref_ps = np.array([fh(x) for x in pg["persite"]])
        assert np.all(np.abs(persite - ref_ps) <= 4e-16 * np.abs(ref_ps) * 4)   # <= 2 ulp of log x weight
        e.update_sumtable(root.clv_index, rb.clv_index, root.scaler_index, rb.scaler_index)
        for t, d1, d2 in pg["derivatives"]:
            a, b = e.derivatives(fh(t))
            assert rel(a, fh(d1)) < 1e-11 and rel(b, fh(d2)) < 1e-11, (fh(t), a, b)
        # all-ones again: the plain path on the compressed columns (a different number)
        e.set_pattern_weights(None)
        plain = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                     rb.scaler_index, root.pmatrix_index)
        e.set_pattern_weights(np.ones(len(w), dtype=np.uint32))
        ones = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                    rb.scaler_index, root.pmatrix_index)
        assert plain == ones and plain != lnl
    finally:
        e.close()


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_optimize_branch_and_nni_batch_match_the_reference(name):
    from pt_b200.engine import ops_array
    pg = gold_of(name)
    tree, seqs, cseqs, w, g, e = compressed_engine(name)
    try:
        for ob in pg["optimize_branch"]:
            node = tree.by_index[ob["node"]]
            ops, midx, mlen = tree.full_traversal(node)
            e.update_prob_matrices(midx, mlen)
            e.update_partials(ops_array(ops))
            nb = node.back
            ln, iters = e.optimize_branch(node.clv_index, node.scaler_index, nb.clv_index,
                                          nb.scaler_index, node.pmatrix_index, fh(ob["len0"]))
            assert abs(ln - fh(ob["len"])) < 1e-9 * max(1.0, fh(ob["len"]))
            assert iters == ob["iters"]
            lnl = e.edge_loglikelihood(node.clv_index, node.scaler_index, nb.clv_index,
                                       nb.scaler_index, node.pmatrix_index)
            assert rel(lnl, fh(ob["lnl"])) < 1e-11
            e.update_prob_matrices([node.pmatrix_index], [fh(ob["len0"])])
    finally:
        e.close()
    tree, seqs, cseqs, w, g, e = compressed_engine(name, directional=True)
    try:
        br = tree.branches()
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(tree.directional_plan()))
        cands = tree.nni_candidates()
        lens, lnls, st = e.score_nni_batch(cand_array(cands))
        assert len(cands) == len(pg["nni"]) == len(g["nni"])
        for (pidx, kind, _, _, _), gl, gu, ln, ll, status in zip(cands, pg["nni"], g["nni"], lens, lnls, st):
            assert (pidx, kind) == (gl["node"], gl["type"]) and status == 0
            assert rel(ll, fh(gl["lnl"])) < 1e-10, (pidx, kind, ll, fh(gl["lnl"]))
            assert abs(ln - fh(gl["len"])) < 1e-8 * max(1.0, fh(gl["len"]))
            assert rel(ll, fh(gu["lnl"])) < 1e-9          # the uncompressed reference run
    finally:
        e.close()


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_whole_tree_and_local_optimisation_on_patterns(name):
    """the step-program kernel with weights: OptimizeAllBranches against the reference on
    the compressed columns, the radius-1 batch against the reference's uncompressed scores"""
    from pt_b200.engine import ops_array
    from pt_b200.local_blo import Program
    pg = gold_of(name)
    tree, seqs, cseqs, w, g, e = compressed_engine(name)
    try:
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        flags = ab.Flags()
        ab.traverse(tree.root, flags, partial=False)
        n_edges = 2 * tree.tip_count - 3
        steps, n_pro, n_body, n_bodies, edge, snaps = ab.program(tree.root, flags, n_edges)
        lengths0 = np.zeros(n_edges)
        for r in tree.by_index.values():
            lengths0[r.pmatrix_index] = r.length
        out, lnl, passes = e.optimize_all_branches(steps, n_pro, n_body, n_bodies, edge, lengths0)
        assert rel(lnl, fh(pg["oab_lnl"])) < 1e-9
        assert rel(lnl, fh(pg["oab_lnl_uncompressed"])) < 1e-9
        for pm, v in pg["oab_lengths"].items():
            assert abs(out[int(pm)] - fh(v)) < 1e-6 * max(1.0, fh(v)), pm
    finally:
        e.close()
    with open(os.path.join(GOLD, "neighborhood_golden.json")) as f:
        gold = json.load(f)[name]["1"]
    tree, seqs, cseqs, w, g, e = compressed_engine(name, directional=True)
    try:
        br = tree.branches()
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(tree.directional_plan()))
        n = tree.tip_count
        prog = Program(tree, 1, n, 3 * (n - 2), 3 * (n - 2), 2 * n - 3)
        lens, lnl, status, sm = e.optimize_local_batch(prog)
        assert prog.count == len(gold)
        for i, row in enumerate(gold):
            assert prog.moves[i] == (row["node"], row["type"]) and status[i] == 0
            assert rel(lnl[i], fh(row["lnl"])) < 1e-9, (i, lnl[i], fh(row["lnl"]))
    finally:
        e.close()


def test_scaler_counts_of_patterns_equal_those_of_their_sites(oracle):
    """300-taxon comb with long branches (scaler counts > 0), 2000 sites drawn from 150
    distinct columns: CLV updates do not look at the weights, so a pattern carries exactly
    the scaler counts of every site it stands for, and the weighted lnL equals the plain lnL
    of the uncompressed alignment (1e-11: another summation order)"""
    from pt_b200.engine import compress_site_patterns, ops_array
    rng = np.random.default_rng(11)
    n, sites, distinct = 300, 2000, 150
    inner = "t2:0.4"
    for i in range(3, n):
        inner = f"({inner},t{i}:0.4):0.4"
    tree = Tree("(t0:0.4,t1:0.4," + inner + ");")
    labels = [f"t{i}" for i in range(n)]
    pool = rng.choice(list("ACGT-"), (n, distinct))
    pick = rng.integers(0, distinct, sites)
    seqs = {k: "".join(pool[i, pick]) for i, k in enumerate(labels)}
    cs, w = compress_site_patterns([seqs[k] for k in labels])
    cseqs = dict(zip(labels, cs))
    assert int(w.sum()) == sites and len(w) <= distinct
    # site -> pattern through the state codes (the compression returns canonical symbols)
    code = {ch: oracle.nt_mask(ch) for ch in set("ACGT-") | set("".join(cs))}
    pat = {tuple(code[cseqs[k][i]] for k in labels): i for i in range(len(w))}
    site_to_pat = np.array([pat[tuple(code[seqs[k][i]] for k in labels)] for i in range(sites)])
    assert np.array_equal(np.bincount(site_to_pat, minlength=len(w)), w)
    _, _, g = load_set("ds1")
    f, s, r = model_of(g)
    full, comp = make_engine(tree, seqs, f, s, r), make_engine(tree, cseqs, f, s, r)
    try:
        comp.set_pattern_weights(w)
        ops, midx, mlen = tree.full_traversal()
        for e in (full, comp):
            e.update_prob_matrices(midx, mlen)
            e.update_partials(ops_array(ops))
        total = 0
        for op in ops:
            sf, sc = full.get_scaler(op[1]), comp.get_scaler(op[1])
            assert np.array_equal(sf, sc[site_to_pat]), op
            total += int(sf.sum())
        assert total > 0
        root, rb = tree.root, tree.root.back
        args = (root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index, root.pmatrix_index)
        assert rel(comp.edge_loglikelihood(*args), full.edge_loglikelihood(*args)) < 1e-11
        # one Newton solve on the root edge: same optimum
        a, ia = full.optimize_branch(*args, root.length)
        b, ib = comp.optimize_branch(*args, root.length)
        assert abs(a - b) < 1e-9 * max(1.0, a)
    finally:
        full.close()
        comp.close()


def test_bad_weights_are_refused():
    from pt_b200.engine import PtGpuError
    tree, seqs, cseqs, w, g, e = compressed_engine("five")
    try:
        bad = w.copy()
        bad[3] = 0
        with pytest.raises(PtGpuError):
            e.set_pattern_weights(bad)
        with pytest.raises(ValueError):
            e.set_pattern_weights(w[:-1])
    finally:
        e.close()


@pytest.mark.parametrize("config", ["config1", "config2"])
def test_ptw_tables_on_compressed_patterns(config, monkeypatch):
    """the whole DS1 walk (BASELINE configs[0] / configs[1]) with PTGPU_COMPRESS_PATTERNS=1:
    the Partition uploads 934 unique columns instead of 1949 sites.  Same good / visited /
    tested topologies and counts as the reference; the printed scores (6 significant digits)
    may differ in the last printed digit where a value sits on a rounding edge."""
    import sys
    sys.path.insert(0, os.path.join(GOLD))
    from make_ptw_golden import run, summarise
    exe = os.path.join(os.path.dirname(GOLD), "..", "pt_b200", "bin", "ptw_batched")
    if not os.path.exists(exe):
        pytest.fail("pt_b200/bin/ptw_batched not built (pt_b200/host/Makefile drivers needs a pt checkout at build time)")
    with open(os.path.join(GOLD, "ptw_golden.json")) as f:
        gold = json.load(f)[config]
    monkeypatch.setenv("PTGPU_COMPRESS_PATTERNS", "1")
    got = summarise(run(exe, gold["args"]))
    assert got["counts"] == gold["counts"]
    for key in ("good_topologies_sha", "visited_topologies_sha", "tested_topologies_sha"):
        assert got[key] == gold[key], key
    a = {l.split("\t", 1)[1]: float(l.split("\t", 1)[0]) for l in got["good"]}
    b = {l.split("\t", 1)[1]: float(l.split("\t", 1)[0]) for l in gold["good"]}
    for topo, v in b.items():
        assert abs(a[topo] - v) <= 2e-6 * abs(v), topo
