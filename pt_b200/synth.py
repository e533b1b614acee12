"""Deterministic synthetic inputs of the BASELINE shapes: random-addition
tree (seed 20260001), DS1's GTR+G4 model, alignment simulated down the tree (seed
20260002) with 2 % gaps.  Shared by bench.py and the GPU tests; data generation
only -- nothing here is on a measured path."""
import numpy as np

from .engine import gamma_rates_mean
from .pytree import Tree, random_tree_newick

# DS1's GTR+G4 model (tests/golden/data/ds1/RAxML_info.DS1:72,75)
FREQS = np.array([0.234, 0.257, 0.280, 0.229])
SUBST = np.array([0.584812, 1.048321, 0.673188, 1.823979, 3.468817, 1.0])
SEED_TREE, SEED_ALN = 20260001, 20260002


DS1_ALPHA = 0.134946


def golden_rates(alpha=DS1_ALPHA):
    """Gamma(alpha) mean rates of 4 categories exactly as the reference computes them
    (pll_compute_gamma_cats, libpll gamma.c:220-292; the engine's host routine is pinned
    bit for bit to the reference's values in tests/test_abi.py)"""
    return gamma_rates_mean(alpha, 4)


RATES = golden_rates()


def gtr_pmats(t):
    """numpy P(t r_c) for data simulation only (not on any measured path)"""
    pi = FREQS
    r = SUBST / SUBST[5]
    q = np.zeros((4, 4))
    k = 0
    for i in range(4):
        for j in range(i + 1, 4):
            q[i, j] = r[k] * pi[j]
            q[j, i] = r[k] * pi[i]
            k += 1
    q -= np.diag(q.sum(1))
    q /= -(pi * np.diag(q)).sum()
    w, v = np.linalg.eig(q)
    vi = np.linalg.inv(v)
    return np.stack([(v * np.exp(w * t * rc)) @ vi for rc in RATES]).real


def simulate(tree, sites, rng):
    """alignment simulated down `tree` under GTR+G4, 2% gaps; returns {label: bytes}"""
    cat = rng.integers(0, 4, sites)
    state = {}
    root = tree.root
    state[id(root)] = rng.choice(4, sites, p=FREQS / FREQS.sum()).astype(np.uint8)
    out = {}
    stack = [(x.back, state[id(root)]) for x in (root, root.next, root.next.next)]
    while stack:
        node, parent = stack.pop()
        cum = np.cumsum(gtr_pmats(node.length), axis=2)
        u = rng.random(sites)
        child = (u[:, None] > cum[cat, parent]).sum(1).astype(np.uint8).clip(0, 3)
        if node.is_tip:
            sym = np.frombuffer(b"ACGT", dtype=np.uint8)[child].copy()
            sym[rng.random(sites) < 0.02] = ord("-")
            out[node.label] = sym
        else:
            stack.append((node.next.back, child))
            stack.append((node.next.next.back, child))
    return out


def parent_tree(taxa, rank):
    """base tree (seed 20260001) moved `rank` random NNIs away: rank r's parent tree"""
    rng = np.random.default_rng(SEED_TREE)
    t = Tree(random_tree_newick(taxa, rng))
    walk = np.random.default_rng(SEED_TREE + 17)
    for _ in range(rank):
        inner = [n for n in t.nodes[t.tip_count:] if not n.back.is_tip]
        t.nni(inner[int(walk.integers(len(inner)))], int(walk.integers(1, 3)))
    return t
