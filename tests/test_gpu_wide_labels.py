"""A topography walk on MORE than 256 taxa (BASELINE configs[4] is 1000 taxa): pt's
label index is a uint8_t (src/label_dictionary.hpp:12), so the reference -- and the drop-in built
from its unchanged headers -- stops with std::length_error before the first tree.  ptw_batched is
built with the 16-bit dictionary of pt_b200/host/ptw/wide_label_dictionary.hpp pre-included; its
keys are checked on the CPU in tests/test_keys_cpu.py, and its tables on DS1 and the stand-ins
against the reference's in tests/test_gpu_dropin.py / test_gpu_standins.py.  Here: a 300-taxon
walk runs, its tables are well formed, the same with and without the visits ahead of time, and
the start tree's score is the engine's own (Python path through the C ABI)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)

TAXA, SITES = 300, 2000
ARGS = ["--lnl-offset", "-0.5", "--optimization-radius", "0"]
WALK_TIMEOUT = 150   # seconds; the walk below is a handful of visits (a few seconds on a B200)


def make_inputs(d):
    """A well-resolved alignment (every branch >= 0.03 substitutions per site, 2000 sites: each NNI
    away from the generating tree costs many lnL units -- a short alignment on 300 taxa has
    thousands of trees within any offset and the walk never ends) and a START tree two NNI moves
    away from the generating one, so that the walk has improving moves to make."""
    from pt_b200 import synth
    from pt_b200.pytree import NNI_LEFT, Tree, random_tree_newick
    rng = np.random.default_rng(300400)
    tree = Tree(random_tree_newick(TAXA, rng, mean_len=0.05))
    for r in tree.by_index.values():
        r.length = max(r.length, 0.03)
    tree = Tree(tree.newick())
    aln = synth.simulate(tree, SITES, np.random.default_rng(300401))
    inner = [n for n in tree.nodes[tree.tip_count:] if not n.back.is_tip]
    for k in (7, len(inner) // 2):
        tree.nni(inner[k], NNI_LEFT)
    start = Tree(tree.newick())
    tre, fa, info = (os.path.join(d, n) for n in ("wide.tre", "wide.fasta", "RAxML_info.wide"))
    with open(tre, "w") as f:
        f.write(start.newick() + "\n")
    with open(fa, "w") as f:
        for label in sorted(aln):
            f.write(">%s\n%s\n" % (label, bytes(aln[label]).decode()))
    with open(os.path.join(HERE, "golden", "data", "ds1", "RAxML_info.DS1")) as src, open(info, "w") as dst:
        dst.write(src.read())
    return start, aln, tre, fa, info


def walk(exe, d, tre, fa, info, tag, env=None):
    out = {k: os.path.join(d, f"{tag}_{k}.tsv") for k in ("good", "visited", "tested")}
    cmd = [exe, "--raxml-info", info, "--threads", "1", "--visited-trees", out["visited"],
           "--tested-trees", out["tested"]] + ARGS + [tre, fa, out["good"]]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=WALK_TIMEOUT,
                           env=dict(os.environ, **(env or {})))
    except subprocess.TimeoutExpired:
        pytest.fail(f"the {TAXA}-taxon walk did not end within {WALK_TIMEOUT} s")
    return r, {k: (sorted(open(p).read().splitlines()) if os.path.exists(p) else None) for k, p in out.items()}


def test_a_300_taxon_walk(tmp_path):
    from test_keys_cpu import splits_of
    from common import make_engine
    from pt_b200 import synth
    from pt_b200.all_branches import visit
    exe = os.path.join(ROOT, "pt_b200", "bin", "ptw_batched")
    if not os.path.exists(exe):
        pytest.fail("pt_b200/bin/ptw_batched not built (pt_b200/host/Makefile drivers needs a pt checkout at build time)")
    d = str(tmp_path)
    tree, aln, tre, fa, info = make_inputs(d)

    r, tables = walk(exe, d, tre, fa, info, "ahead")
    assert r.returncode == 0, r.stderr[-2000:]
    good, visited, tested = tables["good"], tables["visited"], tables["tested"]
    assert len(good) >= 1 and len(visited) >= 2 and len(visited) >= len(good)   # the walk moved
    assert len(tested) >= 2 * (TAXA - 3) - 1          # the start tree's whole NNI neighbourhood was tested
    labels = set(tree.labels())
    topologies = set()
    for line in good + visited:
        lnl, newick = line.split("\t")
        assert np.isfinite(float(lnl)) and float(lnl) < 0
        t_splits = splits_of(newick)
        assert len(t_splits) == TAXA - 3
        topologies.add(frozenset(t_splits))
        assert set(x for s in t_splits for x in s) <= labels
    assert len(topologies) == len(visited)            # every visited tree is another topology

    # the same walk with every visit computed in place: identical tables, scores included
    r0, tables0 = walk(exe, d, tre, fa, info, "inplace", env={"PTGPU_PREFETCH_VISITS": "0"})
    assert r0.returncode == 0, r0.stderr[-2000:]
    assert tables0 == tables

    # the start tree's score against the engine driven from Python (FULL traversal + OptimizeAllBranches
    # from the same lengths; the wanderer's virtual root is another record, so the branches are visited
    # in another order and the optimum agrees to the procedure's own 1e-6-per-pass stopping rule)
    start = frozenset(splits_of(tree.newick()))
    mine = [float(l.split("\t")[0]) for l in visited if frozenset(splits_of(l.split("\t")[1])) == start]
    assert len(mine) == 1
    e = make_engine(tree, aln, synth.FREQS, synth.SUBST, synth.RATES)
    _, lnl, _ = visit(e, tree)
    assert abs(mine[0] - lnl) < 2e-5 * abs(lnl)   # the table prints 6 significant digits

    # the drop-in built from pt's unchanged headers cannot take this input
    ref_like = os.path.join(ROOT, "pt_b200", "bin", "ptw_batched_refkeys")
    if os.path.exists(ref_like):
        rr, _ = walk(ref_like, d, tre, fa, info, "narrow")
        assert rr.returncode != 0 and "index type is not large enough" in rr.stderr
