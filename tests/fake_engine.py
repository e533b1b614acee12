"""TEST-ONLY stand-in for pt_b200.engine.Engine, built from the plain-C oracle.

It exists so that the CPU test-suite (no GPU here) can exercise the site-shard
plumbing of pt_b200/sharded.py -- shard ranges, sequence slicing, the allreduce
hook protocol, lock-step Newton across ranks -- over torch.distributed/gloo with
world_size 2.  It follows the engine's split-phase protocol: local partial sums
in a host buffer -> hook(ptr, count, stream=None) -> scalar update.  Never
imported by the product."""
import ctypes as C

import numpy as np

from oracle.pyoracle import Oracle


class FakeEngine:
    on_cuda = False

    def __init__(self, tips, clv_buffers, sites, rate_cats, prob_matrices, scale_buffers):
        self.O = Oracle()
        self.tips, self.sites, self.cats = tips, sites, rate_cats
        self.clv, self.scal, self.pm = {}, {}, {}
        self.hook = None
        self.offset, self.global_sites = 0, sites

    def set_shard(self, offset, global_sites):
        self.offset, self.global_sites = offset, global_sites

    def set_allreduce(self, fn):
        self.hook = fn

    def _sum(self, values):
        buf = np.ascontiguousarray(values, dtype=np.float64)
        if self.hook is not None:
            rc = self.hook(None, buf.ctypes.data_as(C.c_void_p), buf.size, None)
            assert rc == 0
        return buf

    def set_model(self, freqs, subst, rates):
        self.freqs, self.rates = np.asarray(freqs, float), np.asarray(rates, float)
        self.ev, self.iv, self.el = self.O.eigen_gtr(freqs, subst)

    def set_tip_states(self, tip, seq):
        assert len(seq) == self.sites
        self.clv[tip] = self.O.tip_clv(seq, self.cats)

    def update_prob_matrices(self, idx, lens):
        for i, t in zip(idx, lens):
            self.pm[int(i)] = self.O.pmatrix(self.el, self.ev, self.iv, self.rates, float(t))

    def update_partials(self, ops):
        for op in ops:
            pc, ps, c1, m1, s1, c2, m2, s2 = [int(x) for x in op]
            par, psc = self.O.update_partial(self.sites, self.cats, self.clv[c1], self.clv[c2],
                                             self.pm[m1], self.pm[m2],
                                             self.scal[s1] if s1 >= 0 else None,
                                             self.scal[s2] if s2 >= 0 else None, want_scaler=ps >= 0)
            self.clv[pc] = par
            if ps >= 0:
                self.scal[ps] = psc

    def edge_loglikelihood(self, pclv, pscaler, cclv, cscaler, matrix):
        v = self.O.edge_lnl(self.sites, self.cats, self.clv[pclv], self.scal.get(pscaler),
                            self.clv[cclv], self.scal.get(cscaler), self.pm[matrix], self.freqs)
        return float(self._sum([v])[0])

    def optimize_branch(self, pclv, pscaler, cclv, cscaler, matrix, length0, max_iter=32, eps=1e-6,
                        map_lambda=0.0):
        st = self.O.sumtable(self.sites, self.cats, self.clv[pclv], self.clv[cclv], self.ev,
                             self.iv, self.freqs)
        length, decreasing, iters = length0, False, 0
        for _ in range(max_iter):
            d1, d2 = self._sum(self.O.derivatives(self.sites, self.cats, st, self.el, self.rates,
                                                  length))
            d1 += map_lambda
            iters += 1
            if abs(d1) < eps:
                break
            length = length + d1 / d2 if d2 < 0.0 else length - d1 / d2
            if length < 0.0:
                length = eps
                if decreasing:
                    break
                decreasing = True
        self.update_prob_matrices([matrix], [length])
        return length, iters

    def get_scaler(self, idx):
        return self.scal[idx]

    def close(self):
        pass
