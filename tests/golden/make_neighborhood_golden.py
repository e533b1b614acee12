#!/usr/bin/env python
"""Golden vectors for radius >= 1 NNI scoring, from the UNMODIFIED reference.

Run where /root/reference exists, after `make -f oracle/Makefile.ref`:

    python tests/golden/make_neighborhood_golden.py

For DS1 and `five`, every NNI candidate of the input tree in Wanderer::QueueMoves
order goes through what BranchNeighborhoodOptimizer::EvaluateMove does
(src/move_tester/branch_neighborhood_optimizer.cpp:27-80) at radius 1, 2 and 3:
NNI, InvalidateEdgeClvs, TraversalUpdate(PARTIAL), OptimizeBranchNeighborhood,
TraversalUpdate(PARTIAL), LogLikelihood.  Recorded per candidate: the score and
every branch length the optimiser changed (hex doubles).  The tree is put back
(NNI undone, lengths restored, FULL traversal) between candidates, as the
reference does by working on a clone.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.pyoracle import RefLib  # noqa: E402
from pytree import Tree  # noqa: E402

DATA = os.path.join(HERE, "data")
SETS = {
    "ds1": ("ds1/RAxML_bestTree.DS1", "ds1/DS1.fasta", "ds1/RAxML_info.DS1"),
    "five": ("five/RAxML_bestTree.five", "five/five.fasta",
             "five/RAxML_info.five"),
}
RADII = (1, 2, 3)


def main():
    R = RefLib()
    out = {}
    for name, files in SETS.items():
        tre, fa, info = (os.path.join(DATA, x) for x in files)
        t = Tree(open(tre).readline())
        p = R.open(tre, fa, info)
        p.traversal_update(p.root, True)
        g = out[name] = {}
        for radius in RADII:
            rows = g[str(radius)] = []
            for (pidx, kind, _, _, _) in t.nni_candidates():
                before = {r["pmatrix"]: (i, r["length"]) for i, r in enumerate(p.records())}
                p.nni(pidx, kind)
                p.invalidate_edge(pidx)
                p.traversal_update(pidx, False)
                p.optimize_branch_neighborhood(pidx, radius)
                p.traversal_update(pidx, False)
                lnl = p.loglikelihood(pidx)
                changed = {}
                for i, r in enumerate(p.records()):
                    if r["length"] != before[r["pmatrix"]][1]:
                        changed[str(r["pmatrix"])] = float(r["length"]).hex()
                # undo: topology, lengths, CLVs
                p.nni(pidx, kind)
                for i, r in enumerate(p.records()):
                    if str(r["pmatrix"]) in changed:
                        p.update_branch_length(i, before[r["pmatrix"]][1])
                p.invalidate_edge(pidx)
                p.traversal_update(p.root, True)
                rows.append(dict(node=int(pidx), type=int(kind), lnl=float(lnl).hex(),
                                 lengths=changed))
            print(name, radius, len(rows), min(float.fromhex(r["lnl"]) for r in rows),
                  max(float.fromhex(r["lnl"]) for r in rows))
        p.close()
    with open(os.path.join(HERE, "neighborhood_golden.json"), "w") as fh:
        json.dump(out, fh, indent=0, sort_keys=True)


if __name__ == "__main__":
    main()
