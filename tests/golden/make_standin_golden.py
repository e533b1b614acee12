#!/usr/bin/env python
"""Stand-ins for the hohna data sets the reference repository does not carry (DS2-DS8;
only DS1 is in test/test-data, and there is no network): alignments of the published
SHAPES (taxa x sites), simulated down a seeded random tree under DS1's GTR+G4 model, with
that tree as the start tree and DS1's RAxML_info as the model file.  The UNMODIFIED
reference (oracle/_ref/ptw_threads, --threads 1) walks them and its sorted good / visited /
tested tables become the golden vectors tests/golden/standin_golden.json; the GPU tests run
ptw_batched on the same files.  These are NOT the real DS2-DS8 -- "inputs unavailable on
this box" stands -- they widen the walk's parity to other tree sizes,
including 64 taxa.

    python tests/golden/make_standin_golden.py
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_ptw_golden import digest, score_buckets, summarise  # noqa: E402
from pt_b200.pytree import Tree, random_tree_newick  # noqa: E402
from pt_b200 import synth  # noqa: E402

DATA = os.path.join(HERE, "data", "standins")
# published shapes: name -> (taxa, sites, walk arguments)
CONFIG2 = ["--lnl-offset", "-5", "--optimization-radius", "1"]   # BASELINE configs[1]'s own flags
SETS = {
    "DS2": (29, 2520, CONFIG2),
    "DS3": (36, 1812, ["--lnl-offset", "-2", "--optimization-radius", "1"]),
    "DS4": (41, 1137, CONFIG2),
    "DS5": (50, 378, ["--lnl-offset", "-0.25", "--optimization-radius", "1"]),
    "DS6": (50, 1133, ["--lnl-offset", "-2", "--optimization-radius", "0"]),
    "DS7": (59, 1824, ["--lnl-offset", "-2", "--optimization-radius", "1"]),   # offset -5: > 3 h on one core, given up
    "DS8": (64, 1008, ["--lnl-offset", "-1.5", "--optimization-radius", "0"]),
}


def paths(name):
    d = os.path.join(DATA, name)
    return d, os.path.join(d, name + ".tre"), os.path.join(d, name + ".fasta"), os.path.join(d, "RAxML_info." + name)


def make_inputs(name, taxa, sites):
    d, tre, fa, info = paths(name)
    os.makedirs(d, exist_ok=True)
    seed = 20260100 + int(name[2:])
    rng = np.random.default_rng(seed)
    tree = Tree(random_tree_newick(taxa, rng, mean_len=0.05))
    aln = synth.simulate(tree, sites, np.random.default_rng(seed + 1000))
    with open(tre, "w") as f:
        f.write(tree.newick() + "\n")
    with open(fa, "w") as f:
        for label in sorted(aln):
            f.write(">%s\n%s\n" % (label, bytes(aln[label]).decode()))
    shutil.copy(os.path.join(HERE, "data", "ds1", "RAxML_info.DS1"), info)


def run(binary, name, extra, threads=1, env=None):
    _, tre, fa, info = paths(name)
    with tempfile.TemporaryDirectory() as d:
        out = {k: os.path.join(d, k + ".tsv") for k in ("good", "visited", "tested")}
        cmd = [binary, "--raxml-info", info, "--threads", str(threads), "--visited-trees", out["visited"],
               "--tested-trees", out["tested"]] + extra + [tre, fa, out["good"]]
        subprocess.check_call(cmd, stdout=subprocess.DEVNULL, env=env)
        return {k: sorted(open(p).read().splitlines()) for k, p in out.items()}


if __name__ == "__main__":
    ref = os.path.join(ROOT, "oracle", "_ref", "ptw_threads")
    only = sys.argv[1:]
    # STANDIN_GOLDEN_OUT: another file to write (several sets can be generated side by side and
    # merged afterwards); STANDIN_OFFSET: another --lnl-offset for the sets named on the command line
    path = os.environ.get("STANDIN_GOLDEN_OUT", os.path.join(HERE, "standin_golden.json"))
    if only and os.environ.get("STANDIN_OFFSET"):
        for name in only:
            taxa, sites, extra = SETS[name]
            SETS[name] = (taxa, sites, ["--lnl-offset", os.environ["STANDIN_OFFSET"]] + extra[2:])
    g = json.load(open(path)) if os.path.exists(path) else {}
    for name, (taxa, sites, extra) in SETS.items():
        if only and name not in only:
            continue
        make_inputs(name, taxa, sites)
        t0 = time.time()
        g[name] = dict(taxa=taxa, sites=sites, args=extra, **summarise(run(ref, name, extra)))
        g[name]["score_scale"] = max(abs(float(l.split("\t")[0])) for l in g[name]["good"])
        if len(g[name]["good"]) > 1000:   # keep the fixture small: digest + per-bucket score sums
            g[name]["good_sha"] = digest(g[name]["good"])
            g[name]["good_buckets"] = score_buckets(g[name]["good"])
            del g[name]["good"]
        print(name, taxa, sites, extra, g[name]["counts"], "%.1f s" % (time.time() - t0), flush=True)
        with open(path, "w") as f:
            json.dump(g, f, indent=0)
