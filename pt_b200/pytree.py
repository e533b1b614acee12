"""Small unrooted-tree helper for the tests and the bench driver.

It reproduces libpll's index conventions (tips get clv = pmatrix = 0..n-1 in
parse order, inner nodes clv n.., scaler 0.., three records per inner node;
/root/reference/.../libpll/src/parse_utree.y:250-445) so that operation lists
built here are interchangeable with the reference's, and it builds the
descriptors the C ABI consumes: full-traversal op lists, the directional-CLV
plan (one CLV per inner-node record = per direction) and NNI candidates in
Wanderer::QueueMoves order (/root/reference/src/wanderer.cpp:242-285).
Host-side integer work only; no likelihood arithmetic happens here.
"""
import re

import numpy as np

NNI_LEFT, NNI_RIGHT = 1, 2  # PLL_UTREE_MOVE_NNI_LEFT / _RIGHT (libpll pll.h)


class Rec:
    __slots__ = ("label", "length", "node_index", "clv_index", "scaler_index",
                 "pmatrix_index", "next", "back")

    def __init__(self):
        self.label = None
        self.length = 0.0
        self.node_index = self.clv_index = self.pmatrix_index = 0
        self.scaler_index = -1
        self.next = None
        self.back = None

    @property
    def is_tip(self):
        return self.next is None


def _ring():
    a, b, c = Rec(), Rec(), Rec()
    a.next, b.next, c.next = b, c, a
    return a


_TOKEN = re.compile(r"\s*([(),:;]|[^\s(),:;\[\]]+)")


class Tree:
    def __init__(self, newick):
        toks = _TOKEN.findall(newick.strip())
        self._t, self._i = toks, 0
        self.tip_count = 0
        if self._take() != "(":
            raise ValueError("newick must start with '('")
        subs = [self._subtree()]
        for _ in range(2):
            if self._take() != ",":
                raise ValueError("root must have three children")
            subs.append(self._subtree())
        if self._take() != ")":
            raise ValueError("root must have exactly three children")
        self._label_length()
        if self._take() != ";" or self._i != len(toks):
            raise ValueError("trailing garbage")
        root = _ring()
        for r, s in zip((root, root.next, root.next.next), subs):
            r.back, s.back, r.length = s, r, s.length
        self.root = root
        self._assign_indices()
        self._collect()

    # ---- parsing
    def _peek(self):
        return self._t[self._i] if self._i < len(self._t) else None

    def _take(self):
        t = self._peek()
        self._i += 1
        return t

    def _label_length(self):
        label, length = None, 0.0
        if self._peek() not in (None, "(", ")", ",", ":", ";"):
            label = self._take()
        if self._peek() == ":":
            self._take()
            length = float(self._take())
        return label, length

    def _subtree(self):
        if self._peek() == "(":
            self._take()
            a = self._subtree()
            if self._take() != ",":
                raise ValueError("inner nodes must be binary")
            b = self._subtree()
            if self._take() != ")":
                raise ValueError("inner nodes must be binary")
            n = _ring()
            n.next.back, a.back = a, n.next
            n.next.next.back, b.back = b, n.next.next
            n.label, n.length = self._label_length()
            n.next.length, n.next.next.length = a.length, b.length
            return n
        n = Rec()
        n.label, n.length = self._label_length()
        if n.label is None:
            raise ValueError("tip without label")
        self.tip_count += 1
        return n

    # ---- libpll index template
    def _assign_indices(self):
        cnt = dict(tip=0, clv=self.tip_count, node=self.tip_count, scaler=0)

        def rec(n):
            if n.is_tip:
                n.node_index = n.clv_index = n.pmatrix_index = cnt["tip"]
                n.scaler_index = -1
                cnt["tip"] += 1
                return
            rec(n.next.back)
            rec(n.next.next.back)
            for k, r in enumerate((n, n.next, n.next.next)):
                r.node_index = cnt["node"] + k
                r.clv_index = cnt["clv"]
                r.scaler_index = cnt["scaler"]
            n.pmatrix_index = cnt["clv"]
            n.next.pmatrix_index = n.next.back.pmatrix_index
            n.next.next.pmatrix_index = n.next.next.back.pmatrix_index
            cnt["clv"] += 1
            cnt["scaler"] += 1
            cnt["node"] += 3

        r = self.root
        for x in (r, r.next, r.next.next):
            rec(x.back)
        for k, x in enumerate((r, r.next, r.next.next)):
            x.node_index = cnt["node"] + k
            x.clv_index = cnt["clv"]
            x.scaler_index = cnt["scaler"]
            x.pmatrix_index = x.back.pmatrix_index

    def _collect(self):
        tips, inner = [], []

        def rec(n):
            if n.is_tip:
                tips.append(n)
                return
            rec(n.next.back)
            rec(n.next.next.back)
            inner.append(n)

        r = self.root
        for x in (r, r.next, r.next.next):
            rec(x.back)
        inner.append(r)
        self.nodes = tips + inner
        self.inner_count = len(inner)
        self.by_index = {}
        for n in self.nodes:
            self.by_index[n.node_index] = n
            if n.next:
                self.by_index[n.next.node_index] = n.next
                self.by_index[n.next.next.node_index] = n.next.next

    # ---- descriptors
    def labels(self):
        return [n.label for n in self.nodes[: self.tip_count]]

    def branches(self):
        """{pmatrix_index: length} over all edges"""
        out = {}
        for r in self.by_index.values():
            out[r.pmatrix_index] = r.length
        return out

    def full_traversal(self, root=None):
        """(ops rows, matrix indices, lengths) of a FULL post-order traversal at
        `root` (pll_utree_traverse + pll_utree_create_operations,
        libpll utree.c:284-329,379-442)."""
        root = root or self.root
        order = []

        def post(n):
            if not n.is_tip:
                post(n.next.back)
                post(n.next.next.back)
            order.append(n)

        post(root.back)
        post(root)
        ops, midx, mlen = [], [], []
        for n in order:
            if n is not order[-1].back:
                midx.append(n.pmatrix_index)
                mlen.append(n.length)
            if not n.is_tip:
                a, b = n.next.back, n.next.next.back
                ops.append((n.clv_index, n.scaler_index, a.clv_index, a.pmatrix_index,
                            a.scaler_index, b.clv_index, b.pmatrix_index, b.scaler_index))
        return ops, midx, mlen

    # directional CLVs: record r (inner) owns buffer tips + (r.node_index - tips),
    # holding the CLV of r's node looking away from r.back
    def dir_clv(self, r):
        return r.clv_index if r.is_tip else r.node_index

    def dir_scaler(self, r):
        return -1 if r.is_tip else r.node_index - self.tip_count

    def directional_plan(self, needed=None):
        """ops computing the directional CLVs in dependency order: all 3(n-2) of them, or --
        with `needed` = an iterable of directional CLV indices (dir_clv values) -- only those
        and the ones they are computed from.  An NNI batch never reads a CLV that faces a tip
        (a third of them) nor the one that faces a cherry from its only inner neighbour."""
        done, ops = set(), []
        if needed is not None:
            roots = [self.by_index[i] for i in sorted(set(int(x) for x in needed))
                     if i >= self.tip_count]
        else:
            roots = [r for n in self.nodes[self.tip_count:] for r in (n, n.next, n.next.next)]

        def need(r):
            if r.is_tip or r.node_index in done:
                return
            stack = [(r, False)]
            while stack:
                x, expanded = stack.pop()
                if x.is_tip or x.node_index in done:
                    continue
                a, b = x.next.back, x.next.next.back
                if expanded:
                    done.add(x.node_index)
                    ops.append((self.dir_clv(x), self.dir_scaler(x),
                                self.dir_clv(a), a.pmatrix_index, self.dir_scaler(a),
                                self.dir_clv(b), b.pmatrix_index, self.dir_scaler(b)))
                else:
                    stack.append((x, True))
                    stack.append((b, False))
                    stack.append((a, False))

        for r in roots:
            need(r)
        return ops

    @staticmethod
    def referenced_clvs(cands):
        """directional CLV indices the candidate rows of nni_candidates() read"""
        out = set()
        for (_, _, pop, qop, _) in cands:
            out.update((pop[2], pop[5], qop[2], qop[5]))
        return out

    def nni_candidates(self):
        """[(node_index of p, type, p-op row, q-op row, length0)] in QueueMoves order"""
        out = []
        for p in self.nodes[self.tip_count:]:
            q = p.back
            if q.is_tip:
                continue
            for kind in (NNI_LEFT, NNI_RIGHT):
                x = p.next.back                     # leaves p's side
                w = p.next.next.back                # stays with p
                if kind == NNI_LEFT:
                    y, z = q.next.back, q.next.next.back
                else:
                    y, z = q.next.next.back, q.next.back
                # after the swap p holds (y, w) and q holds (x, z)
                def child(c):
                    return (self.dir_clv(c), c.pmatrix_index, self.dir_scaler(c))
                pop = (0, -1) + child(y) + child(w)
                qop = (0, -1) + child(x) + child(z)
                out.append((p.node_index, kind, pop, qop, p.length))
        return out

    # ---- edits (libpll utree_moves.c:60-109)
    def nni(self, p, kind):
        q = p.back
        t1 = p.next
        t2 = q.next if kind == NNI_LEFT else q.next.next
        a, b = t1.back, t2.back
        for t, s in ((t1, b), (t2, a)):
            t.back, s.back = s, t
            t.length, t.pmatrix_index = s.length, s.pmatrix_index

    def newick(self):
        def rec(n):
            if n.is_tip:
                return f"{n.label}:{n.length!r}"
            return f"({rec(n.next.back)},{rec(n.next.next.back)}):{n.length!r}"
        r = self.root
        return "(" + ",".join(rec(x.back) for x in (r, r.next, r.next.next)) + ");"


def cand_array(cands):
    from pt_b200.engine import CAND_DTYPE, OP_DTYPE
    arr = np.zeros(len(cands), dtype=CAND_DTYPE)
    for i, (_, _, pop, qop, l0) in enumerate(cands):
        for side, row in (("p", pop), ("q", qop)):
            for k, name in enumerate(OP_DTYPE.names):
                arr[side][name][i] = np.int64(row[k]).astype(OP_DTYPE[name])
        arr["length0"][i] = l0
    return arr


def random_tree_newick(n, rng, mean_len=0.05):
    """random-addition topology (each new tip joins a uniformly chosen edge),
    Exp(mean) branch lengths clipped to [1e-4, 1]"""
    def bl():
        return float(min(1.0, max(1e-4, rng.exponential(mean_len))))

    top = [dict(name=f"t{i}", kids=None, len=bl()) for i in range(3)]
    pool = list(top)
    for i in range(3, n):
        x = pool[int(rng.integers(len(pool)))]
        moved = dict(name=x["name"], kids=x["kids"], len=bl())
        tip = dict(name=f"t{i}", kids=None, len=bl())
        x["name"], x["kids"] = None, [moved, tip]
        pool += [moved, tip]

    def fmt(x):
        if x["kids"] is None:
            return f"{x['name']}:{x['len']!r}"
        return "(" + ",".join(fmt(k) for k in x["kids"]) + f"):{x['len']!r}"

    return "(" + ",".join(fmt(x) for x in top) + ");"
