"""Partition::OptimizeAllBranches on pt_b200.pytree trees (host-side integer work only; the
Python mirror of pt_b200/host/src/pll_partition.cpp used by bench.py and the tests).

The orientation protocol of the shared per-node CLV buffer (one `clv_valid` flag per
inner-node record, /root/reference/lib/libptpll/src/pll_util.cpp:45-94), libpll's
post-order traversal with pruning callback (libpll utree.c:284-329) and
pll_utree_create_operations (:379-442), restated on tests/pytree records so that the GPU
tests can drive the procedure both ways through the C ABI:

  * call by call (update_prob_matrices / update_partials / optimize_branch /
    edge_loglikelihood), which is what the reference does, and
  * recorded as ONE step program for ptgpu_optimize_all_branches, which is what
    pt::pll::Partition::OptimizeAllBranches of this repository does.
"""
import numpy as np

MAX_ITER, EPSILON = 32, 1e-6     # PTPLL/pll_partition.hpp:21-22


class Flags:
    def __init__(self):
        self.valid = {}           # node_index of an inner record -> clv_valid (absent: no data yet)

    def orient(self, node, partial):
        """cb_full_traversal / cb_partial_traversal: 1 = recompute this node's CLV"""
        if node.next is None:
            return 1
        ring = (node, node.next, node.next.next)
        if node.node_index not in self.valid:
            for r, v in zip(ring, (1, 0, 0)):
                self.valid[r.node_index] = v
            return 1
        if partial and self.valid[node.node_index]:
            return 0
        for r, v in zip(ring, (1, 0, 0)):
            self.valid[r.node_index] = v
        return 1

    def snapshot(self):
        return dict(self.valid)


def traverse(root, flags, partial=True):
    """(matrix indices, op rows) of TraversalUpdate(root, PARTIAL|FULL)"""
    order = []

    def rec(n):
        if not flags.orient(n, partial):
            return
        if n.next is not None:
            rec(n.next.back)
            rec(n.next.next.back)
        order.append(n)

    rec(root.back)
    rec(root)
    mats, ops = [], []
    for n in order:
        if n is not order[-1].back:
            mats.append(n)
        if n.next is not None:
            a, b = n.next.back, n.next.next.back
            ops.append((n.clv_index, n.scaler_index, a.clv_index, a.pmatrix_index, a.scaler_index,
                        b.clv_index, b.pmatrix_index, b.scaler_index))
    return mats, ops


def postorder_nodes(root):
    out = []

    def rec(n):
        if n.next is not None:
            rec(n.next.back)
            rec(n.next.next.back)
        out.append(n)

    rec(root.back)
    rec(root)
    return out


def set_length(rec, value):
    rec.length = rec.back.length = value


def stepwise(e, root, flags):
    """the reference's call sequence (PTPLL/pll_partition.cpp:490-582); returns (lnl, passes)"""
    from pt_b200.engine import ops_array

    def update(node):
        mats, ops = traverse(node, flags)
        if mats:
            e.update_prob_matrices([m.pmatrix_index for m in mats], [m.length for m in mats])
        if ops:
            e.update_partials(ops_array(ops))

    def lnl():
        b = root.back
        return e.edge_loglikelihood(root.clv_index, root.scaler_index, b.clv_index, b.scaler_index,
                                    root.pmatrix_index)

    def once():
        for node in postorder_nodes(root):
            if node.next is None:
                node = node.back
            update(node)
            b = node.back
            ln, _ = e.optimize_branch(node.clv_index, node.scaler_index, b.clv_index, b.scaler_index,
                                      node.pmatrix_index, node.length, MAX_ITER, EPSILON)
            set_length(node, ln)

    update(root)
    before = lnl()
    once()
    update(root)
    after = lnl()
    passes, i = 1, 0
    while abs(before - after) > EPSILON and i < MAX_ITER:
        before = after
        once()
        update(root)
        after = lnl()
        passes += 1
        i += 1
    return after, passes


def program(root, flags, n_edges):
    """the same procedure recorded for ptgpu_optimize_all_branches:
    (steps array [rows], n_prologue, n_body, n_bodies, edge array, lengths0, flag snapshots)"""
    from pt_b200.local_blo import BLO_MAT, BLO_NOP, BLO_OP, BLO_OPT, EDGE_DTYPE, STEP_DTYPE
    fresh = [False] * n_edges
    cur = []

    def record(node):
        mats, ops = traverse(node, flags)
        for m in mats:
            if not fresh[m.pmatrix_index]:
                fresh[m.pmatrix_index] = True
                cur.append((BLO_MAT, 0, 0, 0, m.pmatrix_index, 0, 0, 0, 0, m.pmatrix_index))
        for op in ops:
            cur.append((BLO_OP, op[0], op[1], op[2], op[3], op[4], op[5], op[6], op[7], 0))

    def one_pass():
        for node in postorder_nodes(root):
            if node.next is None:
                node = node.back
            record(node)
            b = node.back
            cur.append((BLO_OPT, node.clv_index, node.scaler_index, b.clv_index, node.pmatrix_index,
                        b.scaler_index, 0, 0, 0, node.pmatrix_index))
            fresh[node.pmatrix_index] = True
        record(root)

    record(root)
    prologue, cur = cur, []
    bodies, states = [], []
    while len(bodies) < MAX_ITER + 1:
        one_pass()
        bodies.append(cur)
        cur = []
        states.append((flags.snapshot(), list(fresh)))
        if len(states) >= 2 and states[-1] == states[-2]:
            break
    n_body = max(len(b) for b in bodies)
    rows = list(prologue)
    for b in bodies:
        rows += b + [(BLO_NOP,) + (0,) * 9] * (n_body - len(b))
    steps = np.zeros(len(rows), dtype=STEP_DTYPE)
    for i, r in enumerate(rows):
        steps[i] = r
    edge = np.zeros(1, dtype=EDGE_DTYPE)
    b = root.back
    edge[0] = (root.clv_index, root.scaler_index, b.clv_index, b.scaler_index, root.pmatrix_index, 0)
    return steps, len(prologue), n_body, len(bodies), edge, [s[0] for s in states]


def fused(e, tree, flags):
    """OptimizeAllBranches(tree.root) as ONE engine call (ptgpu_optimize_all_branches), the way
    pt::pll::Partition::OptimizeAllBranchesFused issues it; writes the optimised lengths into
    the tree and returns (lnl, passes).  `e` may be an Engine or a SiteShardedEngine with the
    engines' own exchange: then every rank runs the same program on its shard."""
    n_edges = 2 * tree.tip_count - 3
    steps, n_pro, n_body, n_bodies, edge, _ = program(tree.root, flags, n_edges)
    lengths0 = np.zeros(n_edges)
    for r in tree.by_index.values():
        lengths0[r.pmatrix_index] = r.length
    out, lnl, passes = e.optimize_all_branches(steps, n_pro, n_body, n_bodies, edge, lengths0)
    for r in tree.by_index.values():
        r.length = out[r.pmatrix_index]
    return lnl, passes


def visit(e, tree, flags=None, call_by_call=False):
    """What Wanderer::MoveForward does for a tree it visits (src/wanderer.cpp:150-210) from the
    tree's current lengths: FULL TraversalUpdate, LogLikelihood, OptimizeAllBranches (which
    ends with the tree's lnL).  Returns (lnl before, lnl after, passes)."""
    from pt_b200.engine import ops_array
    flags = flags or Flags()
    root = tree.root
    mats, ops = traverse(root, flags, partial=False)
    e.update_prob_matrices([m.pmatrix_index for m in mats], [m.length for m in mats])
    e.update_partials(ops_array(ops))
    rb = root.back
    lnl0 = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                                root.pmatrix_index)
    lnl, passes = stepwise(e, root, flags) if call_by_call else fused(e, tree, flags)
    return lnl0, lnl, passes
