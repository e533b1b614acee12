"""GPU parity tests: every C-ABI entry point of the hot path against the
plain-C oracle (bit-exact for CLVs / scalers / sumtables given identical P
matrices; <= 4 ulp for P matrices because libdevice expm1 is not glibc's;
1e-12 relative for reductions whose summation tree differs from the
reference's sequential sum; north_star asks for 1e-9 relative on lnL) and
against the committed golden vectors produced by the unmodified reference."""
import numpy as np
import pytest

from common import (fh, fha, load_set, make_engine, model_of, oracle_replay, sha, ulp_diff)
from pytree import Tree, cand_array, random_tree_newick

pytestmark = pytest.mark.gpu

REL = 1e-12   # our tolerance for site-sum reductions (north_star: 1e-9)


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.fixture(scope="module", params=["ds1", "five", "newton"])
def setup(request, oracle):
    from pt_b200.engine import ops_array
    name = request.param
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    e = make_engine(tree, seqs, f, s, r)
    ops, midx, mlen = tree.full_traversal()
    e.update_prob_matrices(midx, mlen)
    e.update_partials(ops_array(ops))
    clv, scal, pm, eig = oracle_replay(oracle, tree, seqs, f, s, r, ops)
    yield dict(name=name, tree=tree, seqs=seqs, g=g, model=(f, s, r), e=e, ops=ops,
               clv=clv, scal=scal, pm=pm, eig=eig)
    e.close()


def test_eigen_bit_exact(setup):
    ev, iv, el = setup["e"].eigen()
    g = setup["g"]
    assert np.array_equal(ev, fha(g["evecs"]))
    assert np.array_equal(iv, fha(g["inv_evecs"]))
    assert np.array_equal(el, fha(g["evals"]))


def test_pmatrix_kernel(setup):
    e, g = setup["e"], setup["g"]
    worst = 0
    for i, v in g["pmatrix"].items():
        got = e.get_pmatrix(int(i))
        exp = fha(v)
        assert np.allclose(got, exp, rtol=1e-14, atol=1e-18)
        # rows are 1 + (tiny sums): compare the off-identity part in ulps of 1.0
        worst = max(worst, float(np.abs(got - exp).max()))
    assert worst < 4 * np.finfo(float).eps
    # zero length -> identity (core_pmatrix_avx.c:103-111)
    e.update_prob_matrices([0], [0.0])
    assert np.array_equal(e.get_pmatrix(0), np.tile(np.eye(4).ravel(), g["cats"]))
    e.update_prob_matrices([0], [setup["tree"].branches()[0]])


def test_clv_and_scalers_bit_exact_given_reference_pmatrices(setup, oracle):
    """upload the reference's P matrices, rerun the traversal: CLVs must hash to the
    golden values and scalers must be identical"""
    from pt_b200.engine import ops_array
    e, g = setup["e"], setup["g"]
    saved = {int(i): e.get_pmatrix(int(i)) for i in g["pmatrix"]}
    for i, v in g["pmatrix"].items():
        e.set_pmatrix(int(i), fha(v))
    e.update_partials(ops_array(setup["ops"]))
    for i, h in g["clv_sha"].items():
        assert sha(e.get_clv(int(i))) == h, f"clv {i}"
    for i, nz in g["scalers_nonzero"].items():
        exp = np.zeros(g["sites"], dtype=np.uint32)
        for site, cnt in nz:
            exp[site] = cnt
        assert np.array_equal(e.get_scaler(int(i)), exp)
    # edge lnL / sumtable / derivatives on bit-identical inputs
    tree = setup["tree"]
    root = tree.by_index[g["root"]]
    rb = root.back
    lnl, ps = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                   rb.scaler_index, root.pmatrix_index, persite=True)
    ref_ps = oracle.edge_lnl(g["sites"], g["cats"], setup["clv"][root.clv_index],
                             setup["scal"].get(root.scaler_index), setup["clv"][rb.clv_index],
                             setup["scal"].get(rb.scaler_index), setup["pm"][root.pmatrix_index],
                             setup["model"][0], persite=True)[1]
    assert ulp_diff(ps, ref_ps) <= 2          # device log vs glibc log
    assert rel(lnl, fh(g["lnl"])) < REL
    e.update_sumtable(root.clv_index, rb.clv_index, root.scaler_index, rb.scaler_index)
    assert sha(e.get_sumtable()) == g["sumtable_sha"]
    for t, d1, d2 in g["derivatives"]:
        a, b = e.derivatives(fh(t))
        assert rel(a, fh(d1)) < 1e-11 and rel(b, fh(d2)) < 1e-11, (fh(t), a, fh(d1), b, fh(d2))
    ob = g["optimize_branch"][0]
    ln, iters = e.optimize_branch(root.clv_index, root.scaler_index, rb.clv_index,
                                  rb.scaler_index, root.pmatrix_index, fh(ob["len0"]))
    assert rel(ln, fh(ob["len"])) < 1e-9 and iters == ob["iters"]
    e.update_prob_matrices([root.pmatrix_index], [fh(ob["len0"])])
    for i, v in saved.items():
        e.set_pmatrix(i, v)
    e.update_partials(ops_array(setup["ops"]))


def test_full_path_with_device_pmatrices(setup):
    """everything on the device, nothing injected: lnL within 1e-12 of the reference"""
    e, g, tree = setup["e"], setup["g"], setup["tree"]
    root = tree.by_index[g["root"]]
    rb = root.back
    lnl = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                               root.pmatrix_index)
    assert rel(lnl, fh(g["lnl"])) < REL
    for i, h in g["clv_sha"].items():
        got = e.get_clv(int(i))
        assert np.allclose(got, setup["clv"][int(i)], rtol=1e-13, atol=0)


def test_nni_batch_matches_reference(setup):
    """the batched scorer against SingleBranchOptimizer run by the reference
    (golden 'nni': one entry per candidate in QueueMoves order)"""
    from pt_b200.engine import ops_array
    tree, seqs, g = setup["tree"], setup["seqs"], setup["g"]
    f, s, r = setup["model"]
    e = make_engine(tree, seqs, f, s, r, directional=True)
    try:
        br = tree.branches()
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(tree.directional_plan()))
        cands = tree.nni_candidates()
        lens, lnls, st = e.score_nni_batch(cand_array(cands))
        assert len(cands) == len(g["nni"])
        for (pidx, kind, _, _, _), gl, ln, ll, status in zip(cands, g["nni"], lens, lnls, st):
            assert (pidx, kind) == (gl["node"], gl["type"])
            assert status == 0
            assert rel(ll, fh(gl["lnl"])) < 1e-10, (pidx, kind, ll, fh(gl["lnl"]))
            assert abs(ln - fh(gl["len"])) < 1e-8 * max(1.0, fh(gl["len"]))
        c = e.counters()
        assert c["candidates"] == len(cands)
    finally:
        e.close()


def test_optimize_branch_edges(setup):
    """OptimizeBranch on the golden edges after orienting the CLVs like
    TraversalUpdate(node, PARTIAL) does"""
    from pt_b200.engine import ops_array
    tree, seqs, g = setup["tree"], setup["seqs"], setup["g"]
    f, s, r = setup["model"]
    e = make_engine(tree, seqs, f, s, r)
    try:
        for ob in g["optimize_branch"]:
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


def test_errors_are_loud():
    from pt_b200.engine import Engine, PtGpuError, ops_array
    e = Engine(4, 2, 10, 4, 5, 2)
    try:
        with pytest.raises(PtGpuError):
            e.update_prob_matrices([0], [0.1])            # model not set
        e.set_model([0.25] * 4, [1] * 6, [1, 1, 1, 1])
        with pytest.raises(PtGpuError):
            e.set_tip_states(0, "ACGTACGTAZ")             # illegal state
        with pytest.raises(PtGpuError):
            e.update_prob_matrices([7], [0.1])            # bad index
        with pytest.raises(PtGpuError):
            e.update_partials(ops_array([(1, 0, 0, 0, -1, 1, 1, -1)]))  # writes a tip
        with pytest.raises(PtGpuError):
            e.update_prob_matrices([0], [-1.0])
    finally:
        e.close()
    with pytest.raises(PtGpuError):
        Engine(4, 2, 10, 3, 5, 2)                         # unsupported category count


def test_batched_tip_upload_equals_per_tip_upload():
    """ptgpu_set_tip_states_batch (the loop of Partition::SetTipStates in one call) against
    ptgpu_set_tip_states: contiguous and scattered host buffers, illegal characters"""
    import ctypes
    from pt_b200.engine import Engine, PtGpuError, ops_array
    rng = np.random.default_rng(5)
    tips, sites = 6, 1000
    aln = rng.choice(np.frombuffer(b"ACGTRYKMSWBDHVN-?acgt", dtype=np.uint8), (tips, sites)).copy()
    ops = ops_array([(6, 0, 0, 0, -1, 1, 1, -1), (7, 1, 2, 2, -1, 3, 3, -1), (8, 2, 6, 4, 0, 7, 5, 1)])

    def clvs(upload):
        e = Engine(tips, 4, sites, 4, 9, 4)
        try:
            e.set_model([0.25] * 4, [1] * 6, [0.1, 0.5, 1.0, 2.4])
            upload(e)
            e.update_prob_matrices(list(range(9)), [0.05 * (i + 1) for i in range(9)])
            e.update_partials(ops)
            return e.get_clv(8).copy()
        finally:
            e.close()

    one = clvs(lambda e: [e.set_tip_states(t, bytes(aln[t])) for t in range(tips)])
    order = [3, 0, 5, 1, 4, 2]
    contiguous = clvs(lambda e: e.set_tip_states_batch(range(tips), [aln[t].ctypes.data for t in range(tips)]))
    rows = {t: np.ascontiguousarray(aln[t]) for t in order}          # separate buffers, shuffled order
    scattered = clvs(lambda e: e.set_tip_states_batch(order, [rows[t].ctypes.data for t in order]))
    assert np.array_equal(one, contiguous) and np.array_equal(one, scattered)

    bad = aln.copy()
    bad[4, 777] = ord("Z")
    e = Engine(tips, 4, sites, 4, 9, 4)
    try:
        with pytest.raises(PtGpuError, match="Illegal state code"):
            e.set_tip_states_batch(range(tips), [bad[t].ctypes.data for t in range(tips)])
    finally:
        e.close()


@pytest.mark.parametrize("cats", [1, 2, 8])
def test_other_category_counts(oracle, cats):
    from pt_b200.engine import ops_array
    tree, seqs, g = load_set("five")
    f, s, _ = model_of(g)
    r = oracle.gamma_rates_mean(0.7, cats)
    e = make_engine(tree, seqs, f, s, r)
    try:
        ops, midx, mlen = tree.full_traversal()
        clv, scal, pm, (ev, iv, el) = oracle_replay(oracle, tree, seqs, f, s, r, ops)
        for i, m in pm.items():
            e.set_pmatrix(i, m)
        e.update_partials(ops_array(ops))
        root, rb = tree.root, tree.root.back
        assert np.array_equal(e.get_clv(root.clv_index), clv[root.clv_index])
        sites = len(seqs[tree.labels()[0]])
        exp = oracle.edge_lnl(sites, cats, clv[root.clv_index], scal.get(root.scaler_index),
                              clv[rb.clv_index], scal.get(rb.scaler_index), pm[root.pmatrix_index], f)
        got = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                   rb.scaler_index, root.pmatrix_index)
        assert rel(got, exp) < REL
        e.update_sumtable(root.clv_index, rb.clv_index)
        st = oracle.sumtable(sites, cats, clv[root.clv_index], clv[rb.clv_index], ev, iv, f)
        assert np.array_equal(e.get_sumtable(), st)
        d = e.derivatives(0.1)
        od = oracle.derivatives(sites, cats, st, el, r, 0.1)
        assert rel(d[0], od[0]) < 1e-11 and rel(d[1], od[1]) < 1e-11
    finally:
        e.close()


def test_scaling_deep_tree_bit_exact(oracle):
    """300-taxon comb with long branches: scaler counts > 0 must match the oracle
    exactly and lnL must agree; ragged site count (not a multiple of the tile)"""
    from pt_b200.engine import ops_array
    rng = np.random.default_rng(7)
    n, sites = 300, 333
    inner = "t2:0.4"
    for i in range(3, n):
        inner = f"({inner},t{i}:0.4):0.4"
    tree = Tree("(t0:0.4,t1:0.4," + inner + ");")
    seqs = {f"t{i}": "".join(rng.choice(list("ACGT-"), sites)) for i in range(n)}
    _, _, g = load_set("ds1")
    f, s, r = model_of(g)
    ops, midx, mlen = tree.full_traversal()
    clv, scal, pm, _ = oracle_replay(oracle, tree, seqs, f, s, r, ops)
    e = make_engine(tree, seqs, f, s, r)
    try:
        for i, m in pm.items():
            e.set_pmatrix(i, m)
        e.update_partials(ops_array(ops))
        root, rb = tree.root, tree.root.back
        assert scal[root.scaler_index].max() >= 1
        for op in ops:
            assert np.array_equal(e.get_scaler(op[1]), scal[op[1]]), op
        assert np.array_equal(e.get_clv(root.clv_index), clv[root.clv_index])
        exp = oracle.edge_lnl(sites, 4, clv[root.clv_index], scal.get(root.scaler_index),
                              clv[rb.clv_index], scal.get(rb.scaler_index), pm[root.pmatrix_index], f)
        got = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                   rb.scaler_index, root.pmatrix_index)
        assert rel(got, exp) < REL
    finally:
        e.close()


def test_large_synthetic_roundtrip(oracle):
    """64 taxa x 20k sites: full traversal + lnL against the oracle, and the
    size-independent property lnL(root edge) == lnL(any other edge) after
    re-orienting (pulley principle) to 1e-11 relative"""
    from pt_b200.engine import ops_array
    rng = np.random.default_rng(11)
    n, sites = 64, 20000
    tree = Tree(random_tree_newick(n, rng))
    cols = rng.choice(list("ACGT"), sites)
    seqs = {}
    for lab in tree.labels():
        mut = rng.random(sites) < 0.15
        seqs[lab] = "".join(np.where(mut, rng.choice(list("ACGT"), sites), cols))
    _, _, g = load_set("ds1")
    f, s, r = model_of(g)
    e = make_engine(tree, seqs, f, s, r)
    try:
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        pm = {i: e.get_pmatrix(i) for i in midx}
        e.update_partials(ops_array(ops))
        root, rb = tree.root, tree.root.back
        got = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                   rb.scaler_index, root.pmatrix_index)
        # oracle with the device's P matrices -> CLVs bit-exact, lnL to 1e-12
        cats = 4
        clv, scal = {}, {}
        for tip in tree.nodes[:n]:
            clv[tip.clv_index] = oracle.tip_clv(seqs[tip.label], cats)
        for op in ops:
            pc, ps, c1, m1, s1, c2, m2, s2 = op
            clv[pc], scal[ps] = oracle.update_partial(sites, cats, clv[c1], clv[c2], pm[m1], pm[m2],
                                                      scal.get(s1), scal.get(s2))
        assert np.array_equal(e.get_clv(root.clv_index), clv[root.clv_index])
        exp = oracle.edge_lnl(sites, cats, clv[root.clv_index], scal.get(root.scaler_index),
                              clv[rb.clv_index], scal.get(rb.scaler_index), pm[root.pmatrix_index], f)
        assert rel(got, exp) < REL
        other = tree.nodes[n + 5]
        ops2, midx2, mlen2 = tree.full_traversal(other)
        e.update_partials(ops_array(ops2))
        ob = other.back
        got2 = e.edge_loglikelihood(other.clv_index, other.scaler_index, ob.clv_index,
                                    ob.scaler_index, other.pmatrix_index)
        assert rel(got2, got) < 1e-11
    finally:
        e.close()
