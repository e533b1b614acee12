"""Pattern compression and the oracle's weighted sums on the CPU: the plain-C
oracle and the C ABI's host-only ptgpu_compress_site_patterns against golden vectors the
unmodified reference produced (tests/golden/make_patterns_golden.py), bit for bit."""
import json
import os

import numpy as np
import pytest

from common import GOLD, fh, load_set, model_of, oracle_replay


def gold_of(name):
    with open(os.path.join(GOLD, "patterns_golden.json")) as f:
        return json.load(f)[name]


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_compression_is_the_reference_s(name, oracle):
    from pt_b200.engine import compress_site_patterns
    pg = gold_of(name)
    _, seqs, _ = load_set(name)
    labels = list(seqs.keys())
    full = [seqs[k] for k in labels]
    for fn in (oracle.compress_site_patterns, compress_site_patterns):
        cs, w = fn(full)
        assert [int(x) for x in w] == pg["weights"]
        assert dict(zip(labels, cs)) == pg["sequences"]
        assert int(w.sum()) == pg["sites"] and len(w) == pg["patterns"]
    # idempotent: the unique columns are already sorted and distinct
    cs2, w2 = compress_site_patterns(cs)
    assert cs2 == cs and np.all(w2 == 1)


def test_compression_edge_cases(oracle):
    from pt_b200.engine import PtGpuError, compress_site_patterns
    # one column; all columns equal; case and the ambiguity synonyms fold into one state code
    assert compress_site_patterns(["A", "C"])[1].tolist() == [1]
    cs, w = compress_site_patterns(["AAAA", "CCCC", "----"])
    assert w.tolist() == [4] and [len(x) for x in cs] == [1, 1, 1]
    cs, w = compress_site_patterns(["Aa", "-N", "tU"])
    assert w.tolist() == [2]
    a = compress_site_patterns(["ACGTACGT", "AAAACCCC"])
    b = oracle.compress_site_patterns(["ACGTACGT", "AAAACCCC"])
    assert a[0] == b[0] and np.array_equal(a[1], b[1]) and int(a[1].sum()) == 8
    with pytest.raises(PtGpuError):
        compress_site_patterns(["AC", "A!"])
    with pytest.raises(ValueError):
        oracle.compress_site_patterns(["AC", "A!"])


def test_against_the_reference_itself_when_it_is_built(oracle):
    from oracle.pyoracle import RefLib, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref is not built here")
    rng = np.random.default_rng(5)
    seqs = ["".join(rng.choice(list("ACGT-NRYacgt"), 500)) for _ in range(6)]
    seqs = [s[:40] * 12 + s[:20] for s in seqs]
    a = RefLib().compress_site_patterns(seqs)
    b = oracle.compress_site_patterns(seqs)
    from pt_b200.engine import compress_site_patterns
    c = compress_site_patterns(seqs)
    assert a[0] == b[0] == c[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[1], c[1])


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_oracle_weighted_sums_match_the_reference(name, oracle):
    pg = gold_of(name)
    tree, _, g = load_set(name)
    f, s, r = model_of(g)
    cseqs = pg["sequences"]
    w = np.array(pg["weights"], dtype=np.uint32)
    ops, _, _ = tree.full_traversal()
    clv, scal, pm, (ev, iv, el) = oracle_replay(oracle, tree, cseqs, f, s, r, ops)
    root, rb = tree.root, tree.root.back
    S = len(w)
    lnl, persite = oracle.edge_lnl_w(S, 4, clv[root.clv_index], scal.get(root.scaler_index),
                                     clv[rb.clv_index], scal.get(rb.scaler_index),
                                     pm[root.pmatrix_index], f, w, persite=True)
    assert lnl == fh(pg["lnl"])
    assert [float(x).hex() for x in persite] == pg["persite"]
    assert abs(lnl - fh(pg["lnl_uncompressed"])) < 1e-9 * abs(lnl)
    st = oracle.sumtable(S, 4, clv[root.clv_index], clv[rb.clv_index], ev, iv, f)
    for t, d1, d2 in pg["derivatives"]:
        assert oracle.derivatives_w(S, 4, st, el, r, w, fh(t)) == (fh(d1), fh(d2))
    ob = pg["optimize_branch"][0]
    assert ob["node"] == root.node_index
    ln, iters, status = oracle.optimize_branch_w(S, 4, st, el, r, w, fh(ob["len0"]))
    assert (ln, iters, status) == (fh(ob["len"]), ob["iters"], 0)
    # weights == None is the unweighted function
    assert oracle.edge_lnl_w(S, 4, clv[root.clv_index], scal.get(root.scaler_index), clv[rb.clv_index],
                             scal.get(rb.scaler_index), pm[root.pmatrix_index], f, None) == \
        oracle.edge_lnl(S, 4, clv[root.clv_index], scal.get(root.scaler_index), clv[rb.clv_index],
                        scal.get(rb.scaler_index), pm[root.pmatrix_index], f)
