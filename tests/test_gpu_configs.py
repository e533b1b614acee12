"""GPU parity at the BASELINE stress shapes (configs[3] and configs[4]).

Sizes the oracle finishes in seconds are compared with it directly (bit-exact
CLVs / scalers given identical P matrices, 1e-9 on lnL and lengths); the full
sizes are checked through size-independent properties: the pulley principle
(lnL is the same from every edge), idempotence, and agreement of two different
device code paths (batched scorer vs the single-edge calls)."""
import ctypes

import numpy as np
import pytest

from common import sha, ulp_diff
from pytree import Tree, cand_array, random_tree_newick
from synth import FREQS, RATES, SUBST, simulate

pytestmark = pytest.mark.gpu


def gamma_rates(alpha, cats=4):
    from pt_b200.engine import load_library
    lib = load_library()
    out = np.empty(cats)
    lib.ptgpu_gamma_rates_mean.argtypes = [ctypes.c_double, ctypes.c_uint,
                                           ctypes.POINTER(ctypes.c_double)]
    assert lib.ptgpu_gamma_rates_mean(alpha, cats, out.ctypes.data_as(
        ctypes.POINTER(ctypes.c_double))) == 0
    return out


def build(taxa, sites, seed, mean_len, rates, directional=False):
    from pt_b200.engine import Engine
    rng = np.random.default_rng(seed)
    tree = Tree(random_tree_newick(taxa, rng, mean_len=mean_len))
    aln = simulate(tree, sites, np.random.default_rng(seed + 1))
    nbuf = 3 * (taxa - 2) if directional else taxa - 2
    e = Engine(taxa, nbuf, sites, 4, 2 * taxa - 3, nbuf)
    e.set_model(FREQS, SUBST, rates)
    for tip in tree.nodes[:taxa]:
        e.set_tip_states(tip.clv_index, bytes(aln[tip.label]))
    return tree, aln, e


@pytest.mark.parametrize("alpha,mean_len", [(0.134946, 0.05), (0.05, 0.3)])
def test_config5_deep_tree_scalers_bit_exact_vs_oracle(oracle, alpha, mean_len):
    """1000 taxa (deep traversal, scaler counts >> 1; alpha = 0.05 with long branches drives
    whole categories to underflow): scalers identical, CLVs bit-identical given the oracle's
    P matrices, per-site lnL incl. any -inf pattern as libpll produces it"""
    from pt_b200.engine import ops_array
    taxa, sites = 1000, 2048
    rates = gamma_rates(alpha)
    tree, aln, e = build(taxa, sites, 777, mean_len, rates)
    try:
        ops, midx, mlen = tree.full_traversal()
        ev, iv, el = oracle.eigen_gtr(FREQS, SUBST)
        pm = {i: oracle.pmatrix(el, ev, iv, rates, t) for i, t in tree.branches().items()}
        for i, m in pm.items():
            e.set_pmatrix(i, m)
        e.update_partials(ops_array(ops))
        clv, scal = {}, {}
        for tip in tree.nodes[:taxa]:
            clv[tip.clv_index] = oracle.tip_clv(bytes(aln[tip.label]), 4)
        for op in ops:
            pc, ps, c1, m1, s1, c2, m2, s2 = op
            clv[pc], scal[ps] = oracle.update_partial(sites, 4, clv[c1], clv[c2], pm[m1], pm[m2],
                                                      scal.get(s1), scal.get(s2))
        root, rb = tree.root, tree.root.back
        assert sha(e.get_clv(root.clv_index)) == sha(clv[root.clv_index])
        assert sha(e.get_clv(rb.clv_index)) == sha(clv[rb.clv_index])
        for i in (root.scaler_index, rb.scaler_index, 0, 499, 997):
            if i >= 0:
                assert np.array_equal(e.get_scaler(i), scal[i]), f"scaler {i}"
        deepest = max(int(scal[root.scaler_index].max()),
                      int(scal[rb.scaler_index].max()) if rb.scaler_index >= 0 else 0)
        assert deepest >= (3 if alpha < 0.1 else 1), deepest    # the case really rescales
        lnl, ps = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                       rb.scaler_index, root.pmatrix_index, persite=True)
        ref, rps = oracle.edge_lnl(sites, 4, clv[root.clv_index], scal.get(root.scaler_index),
                                   clv[rb.clv_index], scal.get(rb.scaler_index),
                                   pm[root.pmatrix_index], FREQS, persite=True)
        assert np.array_equal(np.isinf(ps), np.isinf(rps))       # libpll's -inf sites, not "fixed"
        fin = np.isfinite(rps)
        assert ulp_diff(ps[fin], rps[fin]) <= 2
        if np.isfinite(ref):
            assert abs(lnl - ref) <= 1e-12 * abs(ref)
        else:
            assert lnl == ref
    finally:
        e.close()


@pytest.mark.parametrize("alpha,mean_len", [(0.134946, 0.05), (0.05, 0.3)])
def test_config5_full_size_every_scaler_and_lnl_vs_oracle(oracle, alpha, mean_len):
    """BASELINE configs[4] at its FULL size, 1000 taxa x 50,000 sites, compared with the oracle
    directly (~50 M site-updates, seconds on one core): EVERY one of the 998 scaler buffers
    integer-identical, the two root-edge CLVs bit-identical (given the oracle's P matrices), the
    per-site lnL within 2 ulp with libpll's -inf sites reproduced, the sum within 1e-12; also at
    alpha = 0.05 with long branches (scaler counts >> 1)"""
    from pt_b200.engine import ops_array
    taxa, sites = 1000, 50000
    rates = gamma_rates(alpha)
    tree, aln, e = build(taxa, sites, 4242, mean_len, rates)
    try:
        ops, midx, mlen = tree.full_traversal()
        ev, iv, el = oracle.eigen_gtr(FREQS, SUBST)
        pm = {i: oracle.pmatrix(el, ev, iv, rates, t) for i, t in tree.branches().items()}
        for i, m in pm.items():
            e.set_pmatrix(i, m)
        e.update_partials(ops_array(ops))
        tip_of = {tip.clv_index: tip for tip in tree.nodes[:taxa]}
        clv, scal = {}, {}
        root, rb = tree.root, tree.root.back
        keep = {root.clv_index, rb.clv_index}
        deepest, total = 0, 0

        def fetch(c):
            if c in tip_of:
                return oracle.tip_clv(bytes(aln[tip_of[c].label]), 4)
            return clv[c] if c in keep else clv.pop(c)   # every inner CLV is consumed once

        for op in ops:
            pc, ps, c1, m1, s1, c2, m2, s2 = op
            clv[pc], sc = oracle.update_partial(sites, 4, fetch(c1), fetch(c2), pm[m1], pm[m2],
                                                scal.get(s1), scal.get(s2))
            scal[ps] = sc
            got = e.get_scaler(ps)
            assert np.array_equal(got, sc), f"scaler buffer {ps}"
            deepest = max(deepest, int(sc.max()))
            total += int(sc.sum(dtype=np.int64))
            for s_ in (s1, s2):
                if s_ >= 0 and s_ not in (root.scaler_index, rb.scaler_index):
                    scal.pop(s_, None)
        assert len(ops) == taxa - 2
        assert deepest >= (3 if alpha < 0.1 else 1), deepest
        if not rb.is_tip:
            assert sha(e.get_clv(rb.clv_index)) == sha(clv[rb.clv_index])
        assert sha(e.get_clv(root.clv_index)) == sha(clv[root.clv_index])
        lnl, ps = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                       rb.scaler_index, root.pmatrix_index, persite=True)
        rbclv = clv[rb.clv_index] if not rb.is_tip else oracle.tip_clv(bytes(aln[rb.label]), 4)
        ref, rps = oracle.edge_lnl(sites, 4, clv[root.clv_index], scal.get(root.scaler_index), rbclv,
                                   scal.get(rb.scaler_index), pm[root.pmatrix_index], FREQS, persite=True)
        assert np.array_equal(np.isinf(ps), np.isinf(rps))
        fin = np.isfinite(rps)
        assert ulp_diff(ps[fin], rps[fin]) <= 2
        if np.isfinite(ref):
            assert abs(lnl - ref) <= 1e-12 * abs(ref)
        else:
            assert lnl == ref
        print(f"1000 x 50000, alpha {alpha}: scaler sum {total}, deepest count {deepest}, lnL {lnl!r}")
    finally:
        e.close()


def test_config5_full_size_pulley_and_idempotence():
    """1000 taxa x 50,000 sites on the device: lnL agrees from three different edges (pulley
    principle) within 1e-9, and recomputing the traversal reproduces the same bits"""
    from pt_b200.engine import ops_array
    taxa, sites = 1000, 50000
    tree, aln, e = build(taxa, sites, 4242, 0.05, RATES)
    try:
        vals = []
        inner = tree.nodes[taxa:]
        for root in (tree.root, inner[0], inner[len(inner) // 2]):
            ops, midx, mlen = tree.full_traversal(root)
            e.update_prob_matrices(midx, mlen)
            e.update_partials(ops_array(ops))
            rb = root.back
            vals.append(e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                             rb.scaler_index, root.pmatrix_index))
        assert np.isfinite(vals).all()
        assert max(vals) - min(vals) <= 1e-9 * abs(vals[0]), vals
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        a = sha(e.get_clv(tree.root.clv_index)), sha(e.get_scaler(tree.root.scaler_index))
        e.update_partials(ops_array(ops))
        b = sha(e.get_clv(tree.root.clv_index)), sha(e.get_scaler(tree.root.scaler_index))
        assert a == b
        assert e.counters()["site_updates"] == 5 * (taxa - 2) * sites
    finally:
        e.close()


def test_config4_batch_vs_oracle_and_vs_single_edge_path(oracle):
    """64 taxa x 100,000 sites, all 122 NNI candidates in one batch: a sample of candidates
    against the oracle (two CLV updates, sumtable, Newton, lnL at the optimum), and every
    candidate against the device's own single-edge path (update_partials + optimize_branch +
    edge lnL through the P matrix), which shares no reduction code with the batch"""
    from pt_b200.engine import ops_array
    taxa, sites = 64, 100000
    tree, aln, e = build(taxa, sites, 99, 0.05, RATES, directional=True)
    try:
        br = tree.branches()
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        plan = tree.directional_plan()
        e.update_partials(ops_array(plan))
        cands = tree.nni_candidates()
        lens, lnls, st = e.score_nni_batch(cand_array(cands))
        assert len(cands) in (2 * (taxa - 3), 2 * (taxa - 2)) and not st.any() and np.isfinite(lnls).all()
        # determinism: same bits on a second call
        lens2, lnls2, _ = e.score_nni_batch(cand_array(cands))
        assert np.array_equal(lens, lens2) and np.array_equal(lnls, lnls2)

        # --- device single-edge path for every candidate (scratch buffers: reuse two
        # directional slots that no candidate reads: the root record pair is overwritten last)
        spare = [x for x in range(taxa, taxa + 3 * (taxa - 2))]
        used = set()
        for _, _, pop, qop, _ in cands:
            used.update((pop[2], pop[5], qop[2], qop[5]))
        free = [x for x in spare if x not in used]
        assert len(free) >= 2
        pbuf, qbuf = free[0], free[1]
        saved = {b: (e.get_clv(b), e.get_scaler(b - taxa)) for b in (pbuf, qbuf)}
        worst_len = worst_lnl = 0.0
        for k in range(0, len(cands), 7):
            pidx, kind, pop, qop, l0 = cands[k]
            mat = tree.by_index[pidx].pmatrix_index
            ops2 = [(pbuf, pbuf - taxa) + tuple(pop[2:]), (qbuf, qbuf - taxa) + tuple(qop[2:])]
            e.update_partials(ops_array(ops2))
            ln, _ = e.optimize_branch(pbuf, pbuf - taxa, qbuf, qbuf - taxa, mat, l0)
            ll = e.edge_loglikelihood(pbuf, pbuf - taxa, qbuf, qbuf - taxa, mat)
            e.update_prob_matrices([mat], [l0])
            worst_len = max(worst_len, abs(ln - lens[k]) / max(1.0, abs(ln)))
            worst_lnl = max(worst_lnl, abs(ll - lnls[k]) / abs(ll))
        assert worst_len <= 1e-9 and worst_lnl <= 1e-11, (worst_len, worst_lnl)

        # --- oracle on a sample
        ev, iv, el = oracle.eigen_gtr(FREQS, SUBST)
        pm = {i: e.get_pmatrix(i) for i in br}           # the device's own P matrices
        clv, scal = {}, {}
        for tip in tree.nodes[:taxa]:
            clv[tip.clv_index] = oracle.tip_clv(bytes(aln[tip.label]), 4)
        for op in plan:
            pc, ps, c1, m1, s1, c2, m2, s2 = op
            clv[pc], scal[ps] = oracle.update_partial(sites, 4, clv[c1], clv[c2], pm[m1], pm[m2],
                                                      scal.get(s1), scal.get(s2))
        for k in (0, 1, 40, 77, len(cands) - 1):
            pidx, kind, pop, qop, l0 = cands[k]
            P, sp = oracle.update_partial(sites, 4, clv[pop[2]], clv[pop[5]], pm[pop[3]], pm[pop[6]],
                                          scal.get(pop[4]), scal.get(pop[7]))
            Q, sq = oracle.update_partial(sites, 4, clv[qop[2]], clv[qop[5]], pm[qop[3]], pm[qop[6]],
                                          scal.get(qop[4]), scal.get(qop[7]))
            stab = oracle.sumtable(sites, 4, P, Q, ev, iv, FREQS)
            ln, iters, status = oracle.optimize_branch(sites, 4, stab, el, RATES, l0)
            assert status == 0
            ll = oracle.edge_lnl(sites, 4, P, sp, Q, sq, oracle.pmatrix(el, ev, iv, RATES, ln), FREQS)
            assert abs(lens[k] - ln) <= 1e-9 * max(1.0, ln), (k, lens[k], ln)
            assert abs(lnls[k] - ll) <= 1e-9 * abs(ll), (k, lnls[k], ll)
    finally:
        e.close()


def test_few_work_ring_is_bit_identical_to_direct_loads():
    """One branch on a long alignment (64 taxa x 300k sites: 131 chunks of 9 tiles, fewer work
    items than SMs) takes the 6-stage cp.async ring in the Newton sweeps; PTGPU_FEW_WORK_RING=0
    restores direct loads.  Same arithmetic in the same order: OptimizeBranch's length,
    iteration count, the edge lnL and (d1, d2) must agree to the last bit.  (The switch is read
    once per process, hence two child processes.)"""
    import os
    import subprocess
    import sys
    code = r'''
import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from synth import FREQS, SUBST, RATES, parent_tree, simulate
from common import make_engine
from pt_b200.engine import ops_array
tree = parent_tree(64, 0)
aln = simulate(tree, 300000, np.random.default_rng(3))
e = make_engine(tree, {k: bytes(v) for k, v in aln.items()}, FREQS, SUBST, RATES)
ops, midx, mlen = tree.full_traversal()
e.update_prob_matrices(midx, mlen); e.update_partials(ops_array(ops))
r, b = tree.root, tree.root.back
args = (r.clv_index, r.scaler_index, b.clv_index, b.scaler_index, r.pmatrix_index)
e.update_sumtable(r.clv_index, b.clv_index, r.scaler_index, b.scaler_index)
out = [e.derivatives(t) for t in (0.01, 0.1, 1.0)]
ln, it = e.optimize_branch(*args, 0.3)
print(repr((out, ln, it, e.edge_loglikelihood(*args))))
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    got = []
    for flag in ("1", "0"):
        env = dict(os.environ, PTGPU_FEW_WORK_RING=flag)
        r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True,
                           text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        got.append(r.stdout.strip().splitlines()[-1])
    assert got[0] == got[1], got
    assert "nan" not in got[0]
