"""Shared helpers of the test-suite (data sets, golden vectors, oracle replay)."""
import hashlib
import json
import os

import numpy as np

from pytree import Tree

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
DATA = os.path.join(GOLD, "data")
SETS = {
    "ds1": ("ds1/RAxML_bestTree.DS1", "ds1/DS1.fasta", "ds1/RAxML_info.DS1"),
    "five": ("five/RAxML_bestTree.five", "five/five.fasta", "five/RAxML_info.five"),
    "newton": ("tiny/newton.tre", "tiny/newton.fasta", "tiny/RAxML_info.newton"),
}


def fh(x):
    return float.fromhex(x)


def fha(xs):
    return np.array([float.fromhex(x) for x in xs])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def paths(name):
    return tuple(os.path.join(DATA, x) for x in SETS[name])


def golden(name):
    with open(os.path.join(GOLD, f"{name}_golden.json")) as f:
        return json.load(f)


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


def load_set(name):
    """(Tree, {label: seq}, golden dict)"""
    tre, fa, _ = paths(name)
    return Tree(open(tre).readline()), read_fasta(fa), golden(name)


def model_of(g):
    return fha(g["freqs"]), fha(g["subst"]), fha(g["rates"])


def make_engine(tree, seqs, freqs, subst, rates, directional=False, device=0):
    """Engine with libpll's buffer counts (PTPLL/pll_partition.cpp:47-57); with
    directional=True one CLV + scaler buffer per inner-node record."""
    from pt_b200.engine import Engine
    n = tree.tip_count
    sites = len(next(iter(seqs.values())))
    nbuf = 3 * (n - 2) if directional else n - 2
    e = Engine(n, nbuf, sites, len(rates), 2 * n - 3, nbuf, device=device)
    e.set_model(freqs, subst, rates)
    for tip in tree.nodes[:n]:
        e.set_tip_states(tip.clv_index, seqs[tip.label])
    return e


def oracle_replay(O, tree, seqs, freqs, subst, rates, ops, branch_len=None):
    """run an op list through the C oracle; returns (clv dict, scaler dict, pmats, eigen)"""
    cats = len(rates)
    sites = len(next(iter(seqs.values())))
    ev, iv, el = O.eigen_gtr(freqs, subst)
    lens = branch_len or tree.branches()
    pm = {i: O.pmatrix(el, ev, iv, rates, t) for i, t in lens.items()}
    clv, scal = {}, {}
    for tip in tree.nodes[: tree.tip_count]:
        clv[tip.clv_index] = O.tip_clv(seqs[tip.label], cats)
    for op in ops:
        pc, ps, c1, m1, s1, c2, m2, s2 = [int(x) for x in op]
        par, psc = O.update_partial(sites, cats, clv[c1], clv[c2], pm[m1], pm[m2],
                                    scal[s1] if s1 >= 0 else None,
                                    scal[s2] if s2 >= 0 else None, want_scaler=ps >= 0)
        clv[pc] = par
        if ps >= 0:
            scal[ps] = psc
    return clv, scal, pm, (ev, iv, el)


def ulp_diff(a, b):
    """max distance in units of the last place between two float64 arrays"""
    a = np.ascontiguousarray(a, dtype=np.float64).view(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float64).view(np.int64)
    a = np.where(a < 0, np.int64(-2**63) - a, a)
    b = np.where(b < 0, np.int64(-2**63) - b, b)
    return int(np.abs(a - b).max()) if a.size else 0
