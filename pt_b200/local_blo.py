"""Step programs for ptgpu_optimize_local_batch (include/ptgpu.h).

BranchNeighborhoodOptimizer::EvaluateMove
(/root/reference/src/move_tester/branch_neighborhood_optimizer.cpp:27-80) applies an
NNI, re-orients the two CLVs of the central edge, and runs pll-modules'
pllmod_opt_optimize_branch_lengths_local (PLLMOD/optimize/pll_optimize.c:969-1091).
That routine is a recursion (recomp_iterative, :771-934) over the edges within
`radius` of the node; this module unrolls the recursion, for every NNI candidate of a
tree, into a flat list of CLV updates and edge optimisations of one common shape
(positions that do not exist for a candidate, because the recursion met a tip, are
no-ops).  Integer bookkeeping only -- the arithmetic happens in the kernel.

Works on any tree object with libpll-style records (`next`, `back`, `clv_index`,
`scaler_index`, `pmatrix_index`, `node_index`, `length`) such as tests/pytree.Tree;
the C++ twin is pt::pll::Partition::ScoreNniNeighborhood(tree, radius).
"""
import numpy as np

BLO_NOP, BLO_OP, BLO_OPT, BLO_MAT = 0, 1, 2, 3
NNI_LEFT, NNI_RIGHT = 1, 2

STEP_DTYPE = np.dtype([
    ("type", "<i4"), ("a_clv", "<u4"), ("a_scaler", "<i4"),
    ("b_clv", "<u4"), ("b_matrix", "<u4"), ("b_scaler", "<i4"),
    ("c_clv", "<u4"), ("c_matrix", "<u4"), ("c_scaler", "<i4"),
    ("length_slot", "<u4"),
])
EDGE_DTYPE = np.dtype([
    ("parent_clv", "<u4"), ("parent_scaler", "<i4"), ("child_clv", "<u4"),
    ("child_scaler", "<i4"), ("matrix", "<u4"), ("reserved", "<u4"),
])


def template(radius):
    """step types of one smoothing: recomp(tree, radius) then recomp(tree->back, radius-1)"""
    def recomp(r):
        out = [BLO_OPT]
        if r:
            sub = recomp(r - 1)
            out += [BLO_OP] + sub + [BLO_OP] + sub + [BLO_OP]
        return out
    if radius < 1:
        raise ValueError("optimization radius must be positive")
    return recomp(radius) + recomp(radius - 1)


class _Candidate:
    """symbolic program of one candidate: CLV refs are ('g', clv, scaler) for engine
    buffers or ('p', j) for the candidate's j-th private buffer; matrix refs are
    ('g', index) or ('p', slot)."""

    def __init__(self, dir_clv, dir_scaler, slots=None, lengths=None):
        self.dir_clv, self.dir_scaler = dir_clv, dir_scaler
        self.private = {}     # node (clv_index) -> private buffer
        self.slots = slots if slots is not None else {}   # edge (pmatrix_index) -> length slot
        self.lengths = lengths if lengths is not None else []
        self.steps = []

    def clv(self, rec):
        """the CLV of rec's node as the reference's single per-node buffer holds it"""
        if rec.next is None:
            return ("g", rec.clv_index, -1)
        j = self.private.get(rec.clv_index)
        if j is not None:
            return ("p", j)
        return ("g", self.dir_clv(rec), self.dir_scaler(rec))

    def own(self, rec):
        j = self.private.setdefault(rec.clv_index, len(self.private))
        return ("p", j)

    def matrix(self, rec):
        s = self.slots.get(rec.pmatrix_index)
        return ("g", rec.pmatrix_index) if s is None else ("p", s)

    def slot(self, rec):
        s = self.slots.get(rec.pmatrix_index)
        if s is None:
            s = self.slots[rec.pmatrix_index] = len(self.lengths)
            self.lengths.append(rec.length)
        return s

    # update_partials_and_scalers(parent, right, left), pll_optimize.c:741-768
    def op(self, parent, c1, c2):
        b, c = c1.back, c2.back
        self.steps.append((BLO_OP, self.own(parent), self.clv(b), self.matrix(b),
                           self.clv(c), self.matrix(c), 0))

    def opt(self, rec):
        s = self.slot(rec)
        self.steps.append((BLO_OPT, self.clv(rec), self.clv(rec.back), ("p", s), None, None, s))

    def nop(self, n=1):
        self.steps += [(BLO_NOP,)] * n

    def recomp(self, rec, radius):
        size = len(template_recomp(radius))
        if rec is None:
            self.nop(size)
            return
        self.opt(rec)
        if not radius:
            return
        q = rec.next
        if q is None:
            self.nop(size - 1)
            return
        z = q.next
        self.op(q, rec, z)
        self.recomp(q.back, radius - 1)
        self.op(z, q, rec)
        self.recomp(z.back, radius - 1)
        self.op(rec, z, q)


_TEMPLATES = {}


def template_recomp(radius):
    if radius not in _TEMPLATES:
        _TEMPLATES[radius] = ([BLO_OPT] if radius == 0 else
                              [BLO_OPT, BLO_OP] + template_recomp(radius - 1) + [BLO_OP] +
                              template_recomp(radius - 1) + [BLO_OP])
    return _TEMPLATES[radius]


def apply_nni(p, kind):
    """libpll utree_moves.c:60-109 on records (its own inverse)"""
    q = p.back
    t1 = p.next
    t2 = q.next if kind == NNI_LEFT else q.next.next
    a, b = t1.back, t2.back
    for t, s in ((t1, b), (t2, a)):
        t.back, s.back = s, t
        t.length, t.pmatrix_index = s.length, s.pmatrix_index


class Program:
    """arrays for ptgpu_optimize_local_batch"""

    def __init__(self, tree, radius, tips, clv_buffers, scale_buffers, prob_matrices,
                 dir_clv=None, dir_scaler=None):
        dir_clv = dir_clv or tree.dir_clv
        dir_scaler = dir_scaler or tree.dir_scaler
        body = template(radius)
        self.step_types = np.array([BLO_OP, BLO_OP] + body + body, dtype=np.int32)
        self.n_prologue, self.n_body = 2, len(body)
        self.moves = []
        cands = []
        for p in tree.nodes[tree.tip_count:]:
            if p.back.next is None:
                continue
            for kind in (NNI_LEFT, NNI_RIGHT):
                apply_nni(p, kind)
                q = p.back
                # dry run: which edges get a private length / matrix slot
                c = _Candidate(dir_clv, dir_scaler)
                c.recomp(p, radius)
                c.recomp(q, radius - 1)
                c = _Candidate(dir_clv, dir_scaler, c.slots, c.lengths)
                # TraversalUpdate(node, PARTIAL) after InvalidateEdgeClvs: q's node, then p's
                c.op(q, q.next, q.next.next)
                c.op(p, p.next, p.next.next)
                # first smoothing: nodes not yet re-oriented are read from the engine's
                # directional CLVs; later smoothings: from the private copies that the
                # first one left behind (restored with the new lengths)
                for _ in range(2):
                    c.recomp(p, radius)
                    c.recomp(q, radius - 1)
                assert len(c.steps) == len(self.step_types)
                c.smooth = (c.clv(q), c.clv(p), c.matrix(p))
                c.final = (c.clv(p), c.clv(q), c.matrix(p))
                apply_nni(p, kind)
                cands.append(c)
                self.moves.append((p.node_index, kind))
        self.count = len(cands)
        self.n_private_clvs = max((len(c.private) for c in cands), default=1)
        self.n_lengths = max((len(c.lengths) for c in cands), default=1)
        B, NP, NL = self.count, self.n_private_clvs, self.n_lengths
        clv_total = tips + clv_buffers

        def rclv(i, ref):
            if ref[0] == "g":
                return ref[1], ref[2]
            return clv_total + i * NP + ref[1], scale_buffers + i * NP + ref[1]

        def rmat(i, ref):
            return ref[1] if ref[0] == "g" else prob_matrices + i * NL + ref[1]

        self.steps = np.zeros((len(self.step_types), B), dtype=STEP_DTYPE)
        self.smooth_edges = np.zeros(B, dtype=EDGE_DTYPE)
        self.final_edges = np.zeros(B, dtype=EDGE_DTYPE)
        self.lengths0 = np.full((B, NL), 0.1)
        for i, c in enumerate(cands):
            self.lengths0[i, :len(c.lengths)] = c.lengths
            for k, st in enumerate(c.steps):
                row = self.steps[k, i]
                row["type"] = st[0]
                if st[0] == BLO_NOP:
                    continue
                row["a_clv"], row["a_scaler"] = rclv(i, st[1])
                row["b_clv"], row["b_scaler"] = rclv(i, st[2])
                row["b_matrix"] = rmat(i, st[3])
                if st[0] == BLO_OP:
                    row["c_clv"], row["c_scaler"] = rclv(i, st[4])
                    row["c_matrix"] = rmat(i, st[5])
                row["length_slot"] = st[6]
            for arr, (par, chi, mat) in ((self.smooth_edges, c.smooth), (self.final_edges, c.final)):
                arr[i]["parent_clv"], arr[i]["parent_scaler"] = rclv(i, par)
                arr[i]["child_clv"], arr[i]["child_scaler"] = rclv(i, chi)
                arr[i]["matrix"] = rmat(i, mat)
        self.slot_edges = [dict((s, e) for e, s in c.slots.items()) for c in cands]
        # the engine's (directional) CLVs the programs read: what must be up to date before
        # ptgpu_optimize_local_batch -- never one that faces a tip
        self.referenced_clvs = set()
        for c in cands:
            for st in c.steps:
                for ref in st[1:6]:
                    if isinstance(ref, tuple) and ref[0] == "g" and len(ref) == 3 and ref[1] >= tips:
                        self.referenced_clvs.add(ref[1])
            for ref in c.smooth + c.final:
                if ref[0] == "g" and len(ref) == 3 and ref[1] >= tips:
                    self.referenced_clvs.add(ref[1])
