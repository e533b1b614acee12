#!/usr/bin/env python
"""Golden vectors for the pattern-weight path, from the UNMODIFIED reference.

pt never compresses its alignment, but libpll can: this script runs the reference's own
pll_compress_site_patterns on DS1 and five, builds a reference Partition on the unique
columns, hands it the multiplicities with pll_set_pattern_weights and records what the
reference then computes (oracle/_ref/libpllref.so through oracle/refglue/ref_capi.cpp):
the compressed columns and weights themselves, lnL (+ per-pattern), (d1, d2) on a grid of
lengths at the root edge, OptimizeBranch on three edges, the first QueueMoves batch with
SingleBranchOptimizer semantics and OptimizeAllBranches.  While doing so it checks the
plain-C oracle's weighted functions and its compression bit for bit against the same outputs.

    python tests/golden/make_patterns_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.pyoracle import Oracle, RefLib  # noqa: E402
from common import load_set, model_of  # noqa: E402
from make_golden import hx, sha  # noqa: E402


def one_set(name, O, R):
    tree, seqs, gold = load_set(name)
    f, s, r = model_of(gold)
    labels = list(seqs.keys())
    full = [seqs[k] for k in labels]
    cseqs, w = R.compress_site_patterns(full)
    oseqs, ow = O.compress_site_patterns(full)
    assert cseqs == oseqs and np.array_equal(w, ow)
    assert int(w.sum()) == len(full[0])
    g = dict(name=name, sites=len(full[0]), patterns=len(w), weights=[int(x) for x in w],
             sequences=dict(zip(labels, cseqs)))

    newick = tree.newick()
    p = R.open_mem(newick, labels, cseqs, f, s, r)
    p.set_pattern_weights(w)
    recs = p.records()
    S, Cc = p.sites, p.cats
    assert S == len(w)
    ev, iv, el = p.eigen()
    root = p.root
    rb = recs[root]["back"]
    p.traversal_update(root, True)
    lnl, persite = p.loglikelihood(root, True)
    rs, bs = recs[root]["scaler"], recs[rb]["scaler"]
    olnl, ops_ = O.edge_lnl_w(S, Cc, p.clv(recs[root]["clv"]), p.scaler(rs) if rs >= 0 else None,
                              p.clv(recs[rb]["clv"]), p.scaler(bs) if bs >= 0 else None,
                              p.pmatrix(recs[root]["pmatrix"]), f, w, True)
    assert lnl == olnl and np.array_equal(persite, ops_)
    g["lnl"] = float(lnl).hex()
    g["lnl_uncompressed"] = gold["lnl"]
    g["persite"] = hx(persite)

    st = p.sumtable(root)
    g["derivatives"] = []
    for tlen in [1e-6, 1e-4, 0.01, 0.05, 0.1, 0.5, 1.0, 5.0]:
        a = p.derivatives(root, tlen)
        b = O.derivatives_w(S, Cc, st, el, r, w, tlen)
        assert a == b, (tlen, a, b)
        g["derivatives"].append([float(tlen).hex(), float(a[0]).hex(), float(a[1]).hex()])

    g["optimize_branch"] = []
    order = [int(x) for x in p.nodes_order()]
    inner = order[p.tips:]
    for node in [root] + inner[:2]:
        l0 = p.length(node)
        p.traversal_update(node, False)
        nb = recs[node]["back"]
        ost = O.sumtable(S, Cc, p.clv(recs[node]["clv"]), p.clv(recs[nb]["clv"]), ev, iv, f)
        exp_len, iters, status = O.optimize_branch_w(S, Cc, ost, el, r, w, l0)
        got = p.optimize_branch(node)
        assert got == exp_len, (node, got, exp_len)
        lnl_after = p.loglikelihood(node)
        g["optimize_branch"].append(dict(node=int(node), len0=float(l0).hex(), len=float(got).hex(),
                                         iters=iters, lnl=float(lnl_after).hex()))
        p.update_branch_length(node, l0)
    p.close()

    # first QueueMoves batch, SingleBranchOptimizer semantics (src/move_tester/single_branch_optimizer.cpp:23-64)
    p = R.open_mem(newick, labels, cseqs, f, s, r)
    p.set_pattern_weights(w)
    p.traversal_update(p.root, True)
    g["nni"] = []
    for (pidx, kind, _, _, l0) in tree.nni_candidates():
        p.nni(pidx, kind)
        p.invalidate_edge(pidx)
        p.traversal_update(pidx, False)
        ln = p.optimize_branch(pidx)
        sc = p.loglikelihood(pidx)
        p.update_branch_length(pidx, l0)
        p.nni(pidx, kind)
        p.invalidate_edge(pidx)
        g["nni"].append(dict(node=int(pidx), type=int(kind), len=float(ln).hex(), lnl=float(sc).hex()))
    p.traversal_update(p.root, True)
    p.optimize_all_branches(p.root)
    p.traversal_update(p.root, False)
    g["oab_lnl"] = float(p.loglikelihood(p.root)).hex()
    g["oab_lengths"] = {str(rec["pmatrix"]): float(rec["length"]).hex() for rec in p.records()}
    g["oab_lnl_uncompressed"] = gold["oab_lnl"]
    p.close()
    return g


def main():
    O, R = Oracle(), RefLib()
    out = {}
    for name in ("ds1", "five"):
        g = out[name] = one_set(name, O, R)
        print(f"{name}: {g['sites']} sites -> {g['patterns']} patterns, lnl={float.fromhex(g['lnl']):.9f} "
              f"(uncompressed {float.fromhex(g['lnl_uncompressed']):.9f}), oab={float.fromhex(g['oab_lnl']):.9f} "
              f"(uncompressed {float.fromhex(g['oab_lnl_uncompressed']):.9f})")
    path = os.path.join(HERE, "patterns_golden.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=0, sort_keys=True)
    print(path, os.path.getsize(path) // 1024, "KB")


if __name__ == "__main__":
    main()
