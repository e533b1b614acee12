"""CPU tests: the plain-C oracle against the committed golden vectors (which
were produced by the unmodified reference, tests/golden/make_golden.py) and,
when oracle/_ref/libpllref.so is present, against the reference itself."""
import numpy as np
import pytest

from common import (fh, fha, golden, load_set, model_of, oracle_replay, paths, sha)


@pytest.mark.parametrize("name", ["ds1", "five", "newton"])
def test_oracle_matches_golden(oracle, name):
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    ev, iv, el = oracle.eigen_gtr(f, s)
    assert np.array_equal(ev, fha(g["evecs"]))
    assert np.array_equal(iv, fha(g["inv_evecs"]))
    assert np.array_equal(el, fha(g["evals"]))
    ops = g["full_ops"]
    assert [list(o) for o in tree.full_traversal()[0]] == ops
    clv, scal, pm, _ = oracle_replay(oracle, tree, seqs, f, s, r, ops)
    for i, v in g["pmatrix"].items():
        assert np.array_equal(pm[int(i)], fha(v)), f"pmatrix {i}"
    for i, h in g["clv_sha"].items():
        assert sha(clv[int(i)]) == h, f"clv {i}"
    for i, nz in g["scalers_nonzero"].items():
        exp = np.zeros(g["sites"], dtype=np.uint32)
        for site, cnt in nz:
            exp[site] = cnt
        assert np.array_equal(scal[int(i)], exp)
    root = tree.by_index[g["root"]]
    rb = root.back
    lnl, ps = oracle.edge_lnl(g["sites"], g["cats"], clv[root.clv_index], scal.get(root.scaler_index),
                              clv[rb.clv_index], scal.get(rb.scaler_index), pm[root.pmatrix_index],
                              f, persite=True)
    assert lnl == fh(g["lnl"])
    assert sha(ps) == g["persite_sha"]
    st = oracle.sumtable(g["sites"], g["cats"], clv[root.clv_index], clv[rb.clv_index], ev, iv, f)
    assert sha(st) == g["sumtable_sha"]
    for t, d1, d2 in g["derivatives"]:
        assert oracle.derivatives(g["sites"], g["cats"], st, el, r, fh(t)) == (fh(d1), fh(d2))
    ob = g["optimize_branch"][0]
    ln, iters, status = oracle.optimize_branch(g["sites"], g["cats"], st, el, r, fh(ob["len0"]))
    assert (ln, iters, status) == (fh(ob["len"]), ob["iters"], 0)


def test_reference_kats(oracle):
    """known answers printed in the reference's own tests"""
    g = golden("newton")
    assert abs(fh(g["lnl"]) - (-33.387713)) < 1e-6          # lib/libptpll/test/test_ptpll.cpp:38
    g5 = golden("five")
    assert abs(fh(g5["oab_lnl"]) - (-3737.47)) < 0.01        # test/test_ptw.cpp:146
    gd = golden("ds1")
    assert abs(fh(gd["lnl"]) - (-6486.89)) < 0.1             # RAxML_info.DS1:80 (-6486.937579, RAxML's own optimum)


def test_gamma_rates(oracle):
    g = golden("ds1")
    assert np.array_equal(oracle.gamma_rates_mean(0.134946, 4), fha(g["rates"]))
    with pytest.raises(ValueError):
        oracle.gamma_rates_mean(0.01, 4)
    assert oracle.gamma_rates_mean(0.5, 1)[0] == 1.0


def test_nt_map(oracle):
    for ch, m in zip("ACGTacgt-N?RYKMSWBDHVUXO", [1, 2, 4, 8, 1, 2, 4, 8, 15, 15, 15, 5, 10, 12, 3, 6, 9, 14, 13, 11, 7, 8, 15, 15]):
        assert oracle.nt_mask(ch) == m, ch
    assert oracle.nt_mask("Z") == 0 and oracle.nt_mask("!") == 0


def test_scaling_deep_tree(oracle):
    """a 300-taxon comb with long branches underflows: scaler counts > 0 and the
    oracle's lnL stays finite (libpll scaling semantics, core_partials_avx.c:491-527)"""
    from pytree import Tree
    rng = np.random.default_rng(7)
    n, sites = 300, 40
    nw = "(t0:0.4,t1:0.4,"
    inner = "t2:0.4"
    for i in range(3, n):
        inner = f"({inner},t{i}:0.4):0.4"
    nw += inner + ");"
    tree = Tree(nw)
    seqs = {f"t{i}": "".join(rng.choice(list("ACGT"), sites)) for i in range(n)}
    g = golden("ds1")
    f, s, r = model_of(g)
    ops, _, _ = tree.full_traversal()
    clv, scal, pm, _ = oracle_replay(oracle, tree, seqs, f, s, r, ops)
    root = tree.root
    assert scal[root.scaler_index].max() >= 1
    lnl = oracle.edge_lnl(sites, 4, clv[root.clv_index], scal[root.scaler_index],
                          clv[root.back.clv_index], scal.get(root.back.scaler_index),
                          pm[root.pmatrix_index], f)
    assert np.isfinite(lnl) and lnl < -1000


@pytest.mark.parametrize("name", ["ds1", "five", "newton"])
def test_oracle_matches_reference_live(oracle, reflib, name):
    tre, fa, info = paths(name)
    tree, seqs, g = load_set(name)
    p = reflib.open(tre, fa, info)
    try:
        f, s, r = p.model()
        assert np.array_equal(oracle.gamma_rates_mean(
            {"ds1": 0.134946}.get(name, None) or 1.0, 4), r) or name != "ds1"
        ops = p.traversal_update(p.root, True)
        clv, scal, pm, (ev, iv, el) = oracle_replay(oracle, tree, seqs, f, s, r, ops.tolist())
        rev = p.eigen()
        assert np.array_equal(rev[0], ev) and np.array_equal(rev[1], iv) and np.array_equal(rev[2], el)
        for op in ops:
            assert np.array_equal(p.clv(int(op[0])), clv[int(op[0])])
            assert np.array_equal(p.scaler(int(op[1])), scal[int(op[1])])
        root = tree.by_index[p.root]
        lnl = oracle.edge_lnl(p.sites, p.cats, clv[root.clv_index], scal.get(root.scaler_index),
                              clv[root.back.clv_index], scal.get(root.back.scaler_index),
                              pm[root.pmatrix_index], f)
        assert lnl == p.loglikelihood(p.root)
    finally:
        p.close()
