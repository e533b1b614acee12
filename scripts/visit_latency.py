"""Latency of one Partition::OptimizeAllBranches-equivalent call (ptgpu_optimize_all_branches, the
step-program kernel) on DS1 and on the DS8-shape stand-in, through the C ABI (run under gpurun)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import load_set, make_engine, model_of   # noqa: E402
from pt_b200 import all_branches as ab                # noqa: E402
from pt_b200.pytree import Tree                       # noqa: E402

tree, seqs, g = load_set("ds1")
e = make_engine(tree, seqs, *model_of(g))
newick = tree.newick()
for rep in range(3):
    ts = []
    for _ in range(20):
        t = Tree(newick)
        t0 = time.perf_counter()
        lnl0, lnl, passes = ab.visit(e, t)
        ts.append(time.perf_counter() - t0)
    print(f"ds1 visit (FULL traversal + lnL + OptimizeAllBranches, {passes} passes): "
          f"best {1e3 * min(ts):.2f} ms, median {1e3 * sorted(ts)[len(ts) // 2]:.2f} ms, lnL {lnl:.6f}")
e.close()
