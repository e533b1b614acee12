"""Do the small step-program kernels of several engines overlap on one GPU?  T host threads,
one engine each (DS1), every thread runs OptimizeAllBranches K times; prints calls/s per T."""
import os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import all_branches as ab
from common import load_set, make_engine, model_of
from pytree import Tree
from pt_b200.engine import ops_array

K = int(sys.argv[1]) if len(sys.argv) > 1 else 30


def setup():
    tree, seqs, g = load_set("ds1")
    model = model_of(g)
    e = make_engine(tree, seqs, *model)
    ops, midx, mlen = tree.full_traversal()
    e.update_prob_matrices(midx, mlen)
    e.update_partials(ops_array(ops))
    flags = ab.Flags()
    ab.traverse(tree.root, flags, partial=False)
    n_edges = 2 * tree.tip_count - 3
    prog = ab.program(tree.root, flags, n_edges)
    lengths0 = np.zeros(n_edges)
    for r in tree.by_index.values():
        lengths0[r.pmatrix_index] = r.length
    return e, prog, lengths0


def work(e, prog, lengths0, out, i):
    steps, n_pro, n_body, n_bodies, edge, snaps = prog
    for _ in range(K):
        _, lnl, passes = e.optimize_all_branches(steps, n_pro, n_body, n_bodies, edge, lengths0)
    out[i] = (lnl, passes)


for T in (1, 2, 4, 8, 16):
    sets = [setup() for _ in range(T)]
    out = [None] * T
    for s in sets:
        work(*s, out, 0) if False else None
    th = [threading.Thread(target=work, args=(*sets[i], out, i)) for i in range(T)]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    dt = time.perf_counter() - t0
    print(f"threads={T} calls={T*K} wall={dt:.3f}s  {T*K/dt:.1f} calls/s  {1e3*dt/K:.2f} ms per call per thread  lnl={out[0]}", flush=True)
    for s in sets: s[0].close()
