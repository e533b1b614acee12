"""oracle/pyoracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings for the two checkers:

* ``Oracle``  -- the plain-C restatement in oracle/pt_oracle.c (built on demand
  with gcc into oracle/libpt_oracle.so; works on any box with gcc).
* ``RefLib``  -- the UNMODIFIED reference compiled in place by
  oracle/Makefile.ref into oracle/_ref/libpllref.so (exists only after
  ``__graft_entry__.build()`` ran in a container that has /root/reference;
  the built .so travels to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  pt_b200/ never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libpt_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpllref.so")

_dp = C.POINTER(C.c_double)
_up = C.POINTER(C.c_uint)


def _d(a):
    return a.ctypes.data_as(_dp)


def _u(a):
    return None if a is None else a.ctypes.data_as(_up)


def build_oracle(force=False):
    src = os.path.join(HERE, "pt_oracle.c")
    if (not force and os.path.exists(ORACLE_SO)
            and os.path.getmtime(ORACLE_SO) >= os.path.getmtime(src)):
        return ORACLE_SO
    subprocess.check_call(
        ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall",
         "-o", ORACLE_SO, src, "-lm"])
    return ORACLE_SO


class Oracle:
    """numpy front-end of pt_oracle.c (layouts: clv[site][cat][4] etc.)."""

    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.pto_edge_lnl.restype = C.c_double
        L.pto_optimize_branch.restype = C.c_double
        L.pto_nt_mask.restype = C.c_uint

    def eigen_gtr(self, freqs, subst):
        f = np.ascontiguousarray(freqs, dtype=np.float64)
        s = np.ascontiguousarray(subst, dtype=np.float64)
        ev = np.empty(16); iv = np.empty(16); el = np.empty(4)
        self.lib.pto_eigen_gtr(_d(f), _d(s), _d(ev), _d(iv), _d(el))
        return ev, iv, el

    def gamma_rates_mean(self, alpha, cats):
        r = np.empty(cats)
        ok = self.lib.pto_gamma_rates_mean(C.c_double(alpha), C.c_uint(cats), _d(r))
        if not ok:
            raise ValueError("invalid alpha")
        return r

    def tip_clv(self, seq, cats):
        b = seq.encode() if isinstance(seq, str) else bytes(seq)
        clv = np.empty(len(b) * cats * 4)
        if not self.lib.pto_tip_clv(b, C.c_uint(len(b)), C.c_uint(cats), _d(clv)):
            raise ValueError("illegal state")
        return clv

    def nt_mask(self, ch):
        return int(self.lib.pto_nt_mask(C.c_int(ord(ch))))

    def pmatrix(self, evals, evecs, inv_evecs, rates, t):
        rates = np.ascontiguousarray(rates, dtype=np.float64)
        out = np.empty(len(rates) * 16)
        self.lib.pto_pmatrix(_d(evals), _d(evecs), _d(inv_evecs), _d(rates),
                             C.c_uint(len(rates)), C.c_double(t), _d(out))
        return out

    def update_partial(self, sites, cats, left, right, lmat, rmat, lscal, rscal,
                       want_scaler=True):
        parent = np.empty(sites * cats * 4)
        pscal = np.zeros(sites, dtype=np.uint32) if want_scaler else None
        self.lib.pto_update_partial(C.c_uint(sites), C.c_uint(cats), _d(parent),
                                    _u(pscal), _d(left), _d(right), _d(lmat),
                                    _d(rmat), _u(lscal), _u(rscal))
        return parent, pscal

    def edge_lnl(self, sites, cats, clvp, sp, clvc, sc, pmat, freqs, persite=False):
        ps = np.empty(sites) if persite else None
        freqs = np.ascontiguousarray(freqs, dtype=np.float64)
        v = self.lib.pto_edge_lnl(C.c_uint(sites), C.c_uint(cats), _d(clvp), _u(sp),
                                  _d(clvc), _u(sc), _d(pmat), _d(freqs),
                                  None if ps is None else _d(ps))
        return (v, ps) if persite else v

    def sumtable(self, sites, cats, clvp, clvc, evecs, inv_evecs, freqs):
        out = np.empty(sites * cats * 4)
        freqs = np.ascontiguousarray(freqs, dtype=np.float64)
        self.lib.pto_sumtable(C.c_uint(sites), C.c_uint(cats), _d(clvp), _d(clvc),
                              _d(evecs), _d(inv_evecs), _d(freqs), _d(out))
        return out

    def derivatives(self, sites, cats, sumtable, evals, rates, t):
        d1 = C.c_double(); d2 = C.c_double()
        rates = np.ascontiguousarray(rates, dtype=np.float64)
        self.lib.pto_derivatives(C.c_uint(sites), C.c_uint(cats), _d(sumtable),
                                 _d(evals), _d(rates), C.c_double(t),
                                 C.byref(d1), C.byref(d2))
        return d1.value, d2.value

    def optimize_branch(self, sites, cats, sumtable, evals, rates, len0,
                        max_iter=32, eps=1e-6, map_lambda=0.0):
        it = C.c_int(); st = C.c_int()
        rates = np.ascontiguousarray(rates, dtype=np.float64)
        v = self.lib.pto_optimize_branch(C.c_uint(sites), C.c_uint(cats),
                                         _d(sumtable), _d(evals), _d(rates),
                                         C.c_double(len0), C.c_uint(max_iter),
                                         C.c_double(eps), C.c_double(map_lambda),
                                         C.byref(it), C.byref(st))
        return v, it.value, st.value


    # --- pattern weights / compression ------------------------------------
    def edge_lnl_w(self, sites, cats, clvp, sp, clvc, sc, pmat, freqs, weights, persite=False):
        ps = np.empty(sites) if persite else None
        freqs = np.ascontiguousarray(freqs, dtype=np.float64)
        self.lib.pto_edge_lnl_w.restype = C.c_double
        v = self.lib.pto_edge_lnl_w(C.c_uint(sites), C.c_uint(cats), _d(clvp), _u(sp),
                                    _d(clvc), _u(sc), _d(pmat), _d(freqs), _u(weights),
                                    None if ps is None else _d(ps))
        return (v, ps) if persite else v

    def derivatives_w(self, sites, cats, sumtable, evals, rates, weights, t):
        d1 = C.c_double(); d2 = C.c_double()
        rates = np.ascontiguousarray(rates, dtype=np.float64)
        self.lib.pto_derivatives_w(C.c_uint(sites), C.c_uint(cats), _d(sumtable),
                                   _d(evals), _d(rates), _u(weights), C.c_double(t),
                                   C.byref(d1), C.byref(d2))
        return d1.value, d2.value

    def optimize_branch_w(self, sites, cats, sumtable, evals, rates, weights, len0,
                          max_iter=32, eps=1e-6, map_lambda=0.0):
        it = C.c_int(); st = C.c_int()
        rates = np.ascontiguousarray(rates, dtype=np.float64)
        self.lib.pto_optimize_branch_w.restype = C.c_double
        v = self.lib.pto_optimize_branch_w(C.c_uint(sites), C.c_uint(cats),
                                           _d(sumtable), _d(evals), _d(rates), _u(weights),
                                           C.c_double(len0), C.c_uint(max_iter),
                                           C.c_double(eps), C.c_double(map_lambda),
                                           C.byref(it), C.byref(st))
        return v, it.value, st.value

    def compress_site_patterns(self, seqs):
        """(compressed sequences, weights) of a list of equally long strings"""
        return _compress(self.lib.pto_compress_site_patterns, seqs)


def _compress(fn, seqs, ret_is_length=False):
    n = len(seqs[0])
    bufs = [C.create_string_buffer(s.encode() if isinstance(s, str) else bytes(s), n + 1) for s in seqs]
    arr = (C.c_char_p * len(seqs))(*[C.cast(b, C.c_char_p) for b in bufs])
    w = np.zeros(n, dtype=np.uint32)
    if ret_is_length:      # the reference glue: (seqs, count, length, weights) -> new length
        m = fn(arr, C.c_int(len(seqs)), C.c_int(n), w.ctypes.data_as(C.POINTER(C.c_uint)))
        if m < 0:
            raise RuntimeError("pattern compression failed")
    else:                  # (seqs, count, &length, weights) -> status
        ln = C.c_uint(n)
        rc = fn(arr, C.c_uint(len(seqs)), C.byref(ln), w.ctypes.data_as(C.POINTER(C.c_uint)))
        if rc:
            raise ValueError("pattern compression: illegal character (status %d)" % rc)
        m = ln.value
    return [b.raw[:m].decode() for b in bufs], w[:m].copy()


def have_ref():
    return os.path.exists(REF_SO)


class RefLib:
    """The reference itself (pt::pll::Partition) through oracle/refglue/ref_capi.cpp."""

    def __init__(self):
        L = self.lib = C.CDLL(REF_SO)
        L.ref_open.restype = C.c_void_p
        L.ref_open_mem.restype = C.c_void_p
        L.ref_last_error.restype = C.c_char_p
        for f in ("ref_loglikelihood", "ref_optimize_branch", "ref_length"):
            getattr(L, f).restype = C.c_double
        L.pll_compute_gamma_cats.argtypes = [C.c_double, C.c_uint, _dp, C.c_int]

    def gamma_cats_mean(self, alpha, cats):
        r = np.empty(cats)
        self.lib.pll_compute_gamma_cats(alpha, cats, _d(r), 0)  # PLL_GAMMA_RATES_MEAN
        return r

    def compress_site_patterns(self, seqs):
        """pll_compress_site_patterns itself: (compressed sequences, weights)"""
        return _compress(self.lib.ref_compress, seqs, ret_is_length=True)

    def open(self, newick, fasta, info, cats=4, tree_index=0, map_mode=False):
        h = self.lib.ref_open(newick.encode(), fasta.encode(), info.encode(),
                              C.c_int(cats), C.c_int(tree_index), C.c_int(int(map_mode)))
        if not h:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return RefPartition(self.lib, h)

    def open_mem(self, newick, labels, seqs, freqs, subst, rates, map_mode=False):
        n = len(labels)
        la = (C.c_char_p * n)(*[s.encode() for s in labels])
        sa = (C.c_char_p * n)(*[s.encode() if isinstance(s, str) else bytes(s) for s in seqs])
        f = np.ascontiguousarray(freqs, dtype=np.float64)
        s = np.ascontiguousarray(subst, dtype=np.float64)
        r = np.ascontiguousarray(rates, dtype=np.float64)
        h = self.lib.ref_open_mem(newick.encode(), C.c_int(n), la, sa, _d(f), _d(s),
                                  _d(r), C.c_int(len(r)), C.c_int(int(map_mode)))
        if not h:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return RefPartition(self.lib, h)


class RefPartition:
    def __init__(self, lib, h):
        self.lib = lib
        self.h = C.c_void_p(h)
        self.tips = lib.ref_tips(self.h)
        self.sites = lib.ref_sites(self.h)
        self.cats = lib.ref_cats(self.h)

    def close(self):
        if self.h:
            self.lib.ref_close(self.h)
            self.h = None

    def set_pattern_weights(self, weights):
        w = np.ascontiguousarray(weights, dtype=np.uint32)
        assert w.size == self.sites
        self.lib.ref_set_pattern_weights(self.h, w.ctypes.data_as(C.POINTER(C.c_uint)))

    def select_tree(self, i):
        assert self.lib.ref_select_tree(self.h, C.c_int(i))

    @property
    def root(self):
        return self.lib.ref_root_index(self.h)

    def records(self):
        """list of dicts, indexed by node_index"""
        out = []
        for i in range(self.lib.ref_record_count(self.h)):
            nx = C.c_int(); bk = C.c_int(); clv = C.c_uint(); sc = C.c_int()
            pm = C.c_uint(); ln = C.c_double(); lab = C.create_string_buffer(256)
            self.lib.ref_record(self.h, C.c_uint(i), C.byref(nx), C.byref(bk), C.byref(clv),
                                C.byref(sc), C.byref(pm), C.byref(ln), lab, C.c_uint(256))
            out.append(dict(next=nx.value, back=bk.value, clv=clv.value, scaler=sc.value,
                            pmatrix=pm.value, length=ln.value, label=lab.value.decode()))
        return out

    def nodes_order(self):
        out = np.empty(2 * self.tips - 2, dtype=np.uint32)
        self.lib.ref_nodes_order(self.h, _u(out))
        return out

    def model(self):
        f = np.empty(4); s = np.empty(6); r = np.empty(self.cats)
        self.lib.ref_model(self.h, _d(f), _d(s), _d(r))
        return f, s, r

    def eigen(self):
        ev = np.empty(16); iv = np.empty(16); el = np.empty(4)
        self.lib.ref_eigen(self.h, _d(ev), _d(iv), _d(el))
        return ev, iv, el

    def clv(self, idx):
        out = np.empty(self.sites * self.cats * 4)
        self.lib.ref_get_clv(self.h, C.c_uint(idx), _d(out))
        return out

    def scaler(self, idx):
        out = np.empty(self.sites, dtype=np.uint32)
        self.lib.ref_get_scaler(self.h, C.c_uint(idx), _u(out))
        return out

    def pmatrix(self, idx):
        out = np.empty(self.cats * 16)
        self.lib.ref_get_pmatrix(self.h, C.c_uint(idx), _d(out))
        return out

    def traversal_update(self, node, full):
        ops = np.zeros((self.tips, 8), dtype=np.uint32)
        n = self.lib.ref_traversal_update(self.h, C.c_uint(node), C.c_int(int(full)), _u(ops))
        if n < 0:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return ops[:n].view(np.int32).copy()

    def loglikelihood(self, node, persite=False):
        ps = np.empty(self.sites) if persite else None
        v = self.lib.ref_loglikelihood(self.h, C.c_uint(node), None if ps is None else _d(ps))
        return (v, ps) if persite else v

    def update_branch_length(self, node, length):
        self.lib.ref_update_branch_length(self.h, C.c_uint(node), C.c_double(length))

    def optimize_branch(self, node):
        return self.lib.ref_optimize_branch(self.h, C.c_uint(node))

    def optimize_all_branches_once(self, node):
        self.lib.ref_optimize_all_branches_once(self.h, C.c_uint(node))

    def optimize_all_branches(self, node):
        self.lib.ref_optimize_all_branches(self.h, C.c_uint(node))

    def optimize_branch_neighborhood(self, node, radius):
        self.lib.ref_optimize_branch_neighborhood(self.h, C.c_uint(node), C.c_int(radius))

    def sumtable(self, node):
        out = np.empty(self.sites * self.cats * 4)
        self.lib.ref_sumtable(self.h, C.c_uint(node), _d(out))
        return out

    def derivatives(self, node, length):
        d1 = C.c_double(); d2 = C.c_double()
        self.lib.ref_derivatives(self.h, C.c_uint(node), C.c_double(length),
                                 C.byref(d1), C.byref(d2))
        return d1.value, d2.value

    def nni(self, node, kind):
        return self.lib.ref_nni(self.h, C.c_uint(node), C.c_int(kind))

    def invalidate_edge(self, node):
        self.lib.ref_invalidate_edge(self.h, C.c_uint(node))

    def length(self, node):
        return self.lib.ref_length(self.h, C.c_uint(node))
