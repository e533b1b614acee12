#!/usr/bin/env python
"""Generate the committed golden vectors from the UNMODIFIED reference.

Run in a container that has /root/reference, after
`make -f oracle/Makefile.ref` built oracle/_ref/libpllref.so:

    python tests/golden/make_golden.py

For every data set (DS1, five, tiny/newton) it drives pt::pll::Partition
through oracle/refglue/ref_capi.cpp and records: model + Gamma rates,
eigensystem, the tree's libpll records, the FULL-traversal op list, all P
matrices, SHA-256 of every CLV, all scaler vectors, lnL (+ per-site), sumtable
hash and (d1, d2) on a grid of lengths at the root edge, OptimizeBranch on
three edges, the first QueueMoves batch of SingleBranchOptimizer results and an
OptimizeAllBranches trace.  Doubles are stored as hex strings (bit exact).
While doing so it checks the plain-C oracle (oracle/pt_oracle.c) bit for bit
against the same reference outputs and aborts on any difference.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.pyoracle import Oracle, RefLib  # noqa: E402
from pytree import Tree, NNI_LEFT, NNI_RIGHT  # noqa: E402

DATA = os.path.join(HERE, "data")
SETS = {
    "ds1": ("ds1/RAxML_bestTree.DS1", "ds1/DS1.fasta", "ds1/RAxML_info.DS1"),
    "five": ("five/RAxML_bestTree.five", "five/five.fasta", "five/RAxML_info.five"),
    "newton": ("tiny/newton.tre", "tiny/newton.fasta", "tiny/RAxML_info.newton"),
}


def hx(a):
    return [float(x).hex() for x in np.atleast_1d(a)]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def read_fasta(path):
    seqs, name = {}, None
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            name = line[1:]
            seqs[name] = ""
        elif name:
            seqs[name] += line
    return seqs


def one_set(name, tre, fa, info, O, R):
    tre, fa, info = (os.path.join(DATA, x) for x in (tre, fa, info))
    p = R.open(tre, fa, info)
    g = dict(name=name, tips=p.tips, sites=p.sites, cats=p.cats)
    f, s, r = p.model()
    ev, iv, el = p.eigen()
    oev, oiv, oel = O.eigen_gtr(f, s)
    assert np.array_equal(ev, oev) and np.array_equal(iv, oiv) and np.array_equal(el, oel)
    g.update(freqs=hx(f), subst=hx(s), rates=hx(r), evecs=hx(ev), inv_evecs=hx(iv), evals=hx(el))
    recs = p.records()
    g["records"] = recs
    g["nodes_order"] = [int(x) for x in p.nodes_order()]
    g["root"] = int(p.root)
    t = Tree(open(tre).readline())
    assert [n.node_index for n in t.nodes] == g["nodes_order"]

    ops = p.traversal_update(p.root, True)
    g["full_ops"] = ops.tolist()
    S, Cc = p.sites, p.cats
    lengths = {rec["pmatrix"]: rec["length"] for rec in recs}
    pms = {}
    for i in range(2 * p.tips - 3):
        pms[i] = p.pmatrix(i)
        assert np.array_equal(pms[i], O.pmatrix(el, ev, iv, r, lengths[i])), i
    g["pmatrix"] = {str(i): hx(v) for i, v in pms.items()}

    seqs = read_fasta(fa)
    clv, scal = {}, {}
    for i in range(p.tips):
        clv[i] = O.tip_clv(seqs[recs[i]["label"]], Cc)
        assert np.array_equal(clv[i], p.clv(i))
    g["clv_sha"], g["scalers_nonzero"] = {}, {}
    for op in ops:
        pc, ps, c1, m1, s1, c2, m2, s2 = [int(x) for x in op]
        par, psc = O.update_partial(S, Cc, clv[c1], clv[c2], pms[m1], pms[m2],
                                    scal.get(s1) if s1 >= 0 else None,
                                    scal.get(s2) if s2 >= 0 else None)
        assert np.array_equal(par, p.clv(pc)) and np.array_equal(psc, p.scaler(ps)), pc
        clv[pc], scal[ps] = par, psc
        g["clv_sha"][str(pc)] = sha(par)
        nz = np.nonzero(psc)[0]
        g["scalers_nonzero"][str(ps)] = [[int(i), int(psc[i])] for i in nz]

    root = p.root
    rb = recs[root]["back"]
    lnl, persite = p.loglikelihood(root, True)
    rs = recs[root]["scaler"]
    bs = recs[rb]["scaler"]
    olnl, ops_ = O.edge_lnl(S, Cc, clv[recs[root]["clv"]], scal.get(rs), clv[recs[rb]["clv"]],
                            scal.get(bs) if bs >= 0 else None, pms[recs[root]["pmatrix"]], f, True)
    assert lnl == olnl and np.array_equal(persite, ops_)
    g["lnl"] = float(lnl).hex()
    g["persite_sha"] = sha(persite)
    g["persite_head"] = hx(persite[:8])

    st = p.sumtable(root)
    ost = O.sumtable(S, Cc, clv[recs[root]["clv"]], clv[recs[rb]["clv"]], ev, iv, f)
    assert np.array_equal(st, ost)
    g["sumtable_sha"] = sha(st)
    g["derivatives"] = []
    for tlen in [1e-6, 1e-4, 0.01, 0.05, 0.1, 0.5, 1.0, 5.0]:
        a = p.derivatives(root, tlen)
        b = O.derivatives(S, Cc, ost, el, r, tlen)
        assert a == b, (tlen, a, b)
        g["derivatives"].append([float(tlen).hex(), float(a[0]).hex(), float(a[1]).hex()])

    # OptimizeBranch on three edges (after PARTIAL traversal towards each)
    g["optimize_branch"] = []
    inner = [n for n in g["nodes_order"][p.tips:]]
    for node in [root] + inner[:2]:
        l0 = p.length(node)
        p.traversal_update(node, False)
        nb = recs[node]["back"]
        ost = O.sumtable(S, Cc, p.clv(recs[node]["clv"]), p.clv(recs[nb]["clv"]), ev, iv, f)
        exp_len, iters, status = O.optimize_branch(S, Cc, ost, el, r, l0)
        got = p.optimize_branch(node)
        assert got == exp_len, (node, got, exp_len)
        lnl_after = p.loglikelihood(node)
        g["optimize_branch"].append(dict(node=int(node), len0=float(l0).hex(),
                                         len=float(got).hex(), iters=iters,
                                         lnl=float(lnl_after).hex()))
        p.update_branch_length(node, l0)  # restore

    p.close()

    # first QueueMoves batch with SingleBranchOptimizer semantics
    # (src/move_tester/single_branch_optimizer.cpp:23-64), fresh partition
    p = R.open(tre, fa, info)
    p.traversal_update(p.root, True)
    g["nni"] = []
    for (pidx, kind, _, _, l0) in t.nni_candidates():
        p.nni(pidx, kind)
        p.invalidate_edge(pidx)
        p.traversal_update(pidx, False)
        ln = p.optimize_branch(pidx)
        sc = p.loglikelihood(pidx)
        p.update_branch_length(pidx, l0)
        p.nni(pidx, kind)
        p.invalidate_edge(pidx)
        g["nni"].append(dict(node=int(pidx), type=int(kind), len=float(ln).hex(),
                             lnl=float(sc).hex()))
    # OptimizeAllBranches from the input lengths
    p.traversal_update(p.root, True)
    p.optimize_all_branches(p.root)
    p.traversal_update(p.root, False)
    g["oab_lnl"] = float(p.loglikelihood(p.root)).hex()
    g["oab_lengths"] = {str(rec["pmatrix"]): float(rec["length"]).hex() for rec in p.records()}
    p.close()
    return g


def main():
    O, R = Oracle(), RefLib()
    for name, (tre, fa, info) in SETS.items():
        g = one_set(name, tre, fa, info, O, R)
        out = os.path.join(HERE, f"{name}_golden.json")
        with open(out, "w") as fh:
            json.dump(g, fh, indent=0, sort_keys=True)
        print(f"{name}: lnl={float.fromhex(g['lnl']):.6f} nni={len(g['nni'])} "
              f"oab_lnl={float.fromhex(g['oab_lnl']):.6f} -> {out} "
              f"({os.path.getsize(out)//1024} KB)")


if __name__ == "__main__":
    main()
