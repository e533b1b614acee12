"""GPU tests of site-sharded operation.

* two shards of DS1 as two engines on ONE GPU, driven by two host threads whose
  allreduce hook is an in-process barrier + sum (no kernel waits on another
  kernel; only the host threads meet) -- runs on the 1-GPU test box;
* the real thing, two processes / two GPUs / NCCL, when the box has >= 2 GPUs.
Both compare against the unsharded engine and the reference's golden values."""
import os
import socket
import subprocess
import sys
import threading

import numpy as np
import pytest

from common import fh, load_set, make_engine, model_of
from pytree import cand_array

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class ThreadAllreduce:
    """sum over the shards of this process: every shard's thread brings its buffer"""

    def __init__(self, n):
        import torch
        self.torch, self.n = torch, n
        self.barrier = threading.Barrier(n)
        self.slots = [None] * n
        self.calls = 0

    def hook_for(self, k):
        from pt_b200.engine import ALLREDUCE_FN
        from pt_b200.sharded import _DevBuf
        torch = self.torch

        def fn(ctx, ptr, count, stream):
            try:
                dev = torch.device("cuda", 0)
                t = torch.as_tensor(_DevBuf(ptr, int(count)), device=dev)
                torch.cuda.ExternalStream(stream, device=dev).synchronize()
                self.slots[k] = t
                self.barrier.wait()
                if k == 0:
                    total = self.slots[0].clone()
                    for other in self.slots[1:]:
                        total += other
                    for s in self.slots:
                        s.copy_(total)
                    torch.cuda.synchronize()
                    self.calls += 1
                self.barrier.wait()
                return 0
            except Exception as exc:  # pragma: no cover
                print("hook failed:", exc, file=sys.stderr)
                self.barrier.abort()
                return 1
        return ALLREDUCE_FN(fn)


@pytest.mark.parametrize("name", ["ds1", "five"])
def test_two_site_shards_on_one_gpu_match_unsharded(name):
    from pt_b200.engine import Engine, ops_array
    from pt_b200.sharded import shard_range
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    n, S = tree.tip_count, g["sites"]
    whole = make_engine(tree, seqs, f, s, r, directional=True)
    br = tree.branches()
    plan = ops_array(tree.directional_plan())
    ops, midx, mlen = tree.full_traversal()
    cands = cand_array(tree.nni_candidates())
    root = tree.by_index[g["root"]]
    rb = root.back

    def run(e, out):
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(ops))
        out["lnl"] = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                          rb.scaler_index, root.pmatrix_index)
        out["opt"] = e.optimize_branch(root.clv_index, root.scaler_index, rb.clv_index,
                                       rb.scaler_index, root.pmatrix_index, root.length)
        out["bounded"] = e.optimize_branch_bounded(root.clv_index, root.scaler_index, rb.clv_index,
                                                   rb.scaler_index, root.length)
        e.update_prob_matrices([root.pmatrix_index], [root.length])
        e.update_partials(plan)
        out["nni"] = e.score_nni_batch(cands)
        out["scalers"] = {i: e.get_scaler(i) for i in range(n - 2)}

    ref = {}
    run(whole, ref)
    whole.close()
    assert abs(ref["lnl"] - fh(g["lnl"])) <= 1e-12 * abs(ref["lnl"])

    world = 2
    ar = ThreadAllreduce(world)
    engines, outs, hooks, errs = [], [{} for _ in range(world)], [], []
    for k in range(world):
        b, e_ = shard_range(S, world, k)
        e = Engine(n, 3 * (n - 2), e_ - b, len(r), 2 * n - 3, 3 * (n - 2))
        e.set_model(f, s, r)
        for tip in tree.nodes[:n]:
            e.set_tip_states(tip.clv_index, seqs[tip.label][b:e_])
        e.set_shard(b, S)
        hooks.append(ar.hook_for(k))
        e.set_allreduce(hooks[-1])
        engines.append(e)

    def worker(k):
        try:
            run(engines[k], outs[k])
        except Exception as exc:
            errs.append(exc)
            ar.barrier.abort()

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errs, errs
    for k in range(world):
        o = outs[k]
        assert abs(o["lnl"] - ref["lnl"]) <= 1e-12 * abs(ref["lnl"])
        assert abs(o["opt"][0] - ref["opt"][0]) <= 1e-10 * max(1.0, ref["opt"][0])
        assert o["opt"][1] == ref["opt"][1]                       # same number of Newton steps
        assert abs(o["bounded"][0] - ref["bounded"][0]) <= 1e-10 and o["bounded"][1:] == ref["bounded"][1:]
        assert np.allclose(o["nni"][0], ref["nni"][0], rtol=1e-9, atol=1e-12)
        assert np.allclose(o["nni"][1], ref["nni"][1], rtol=1e-12, atol=0)
        assert not o["nni"][2].any()
    # both shards saw identical (global) results
    assert outs[0]["lnl"] == outs[1]["lnl"] and outs[0]["opt"] == outs[1]["opt"]
    assert np.array_equal(outs[0]["nni"][1], outs[1]["nni"][1])
    # scaler counts: the shards' rows concatenate to the unsharded rows (integer-exact)
    for i in range(n - 2):
        assert np.array_equal(np.concatenate([outs[k]["scalers"][i] for k in range(world)]),
                              ref["scalers"][i])
    for e in engines:
        e.set_allreduce(None)
        e.close()


NCCL_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import torch, torch.distributed as dist
from common import fh, load_set, model_of
from pt_b200.engine import ops_array
from pt_b200.sharded import SiteShardedEngine
from pytree import cand_array
rank = int(sys.argv[1]); world = int(sys.argv[2])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                        device_id=torch.device("cuda", rank))
for name in ("ds1", "five"):
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    n = tree.tip_count
    e = SiteShardedEngine(n, 3 * (n - 2), g["sites"], len(r), 2 * n - 3, 3 * (n - 2), device=rank)
    e.set_model(f, s, r)
    for tip in tree.nodes[:n]:
        e.set_tip_states(tip.clv_index, seqs[tip.label])
    ops, midx, mlen = tree.full_traversal()
    e.update_prob_matrices(midx, mlen)
    e.update_partials(ops_array(ops))
    root = tree.by_index[g["root"]]; rb = root.back
    lnl = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                               root.pmatrix_index)
    assert abs(lnl - fh(g["lnl"])) <= 1e-12 * abs(lnl), (lnl, fh(g["lnl"]))
    for i, nz in g["scalers_nonzero"].items():
        exp = np.zeros(g["sites"], dtype=np.uint32)
        for site, cnt in nz:
            exp[site] = cnt
        assert np.array_equal(e.gather_scaler(int(i)), exp)
    ob = g["optimize_branch"][0]
    ln, iters = e.optimize_branch(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                                  root.pmatrix_index, fh(ob["len0"]))
    assert abs(ln - fh(ob["len"])) <= 1e-9 * max(1.0, ln) and iters == ob["iters"], (ln, iters)
    br = tree.branches()
    e.update_prob_matrices(list(br.keys()), list(br.values()))
    e.update_partials(ops_array(tree.directional_plan()))
    cands = tree.nni_candidates()
    lens, lnls, st = e.score_nni_batch(cand_array(cands))
    for gl, ln_, ll in zip(g["nni"], lens, lnls):
        assert abs(ll - fh(gl["lnl"])) <= 1e-10 * abs(ll), (ll, fh(gl["lnl"]))
        assert abs(ln_ - fh(gl["len"])) <= 1e-8 * max(1.0, fh(gl["len"]))
    assert not st.any()
    e.close()
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_site_shards_two_gpus_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "worker.py"
    script.write_text(NCCL_WORKER.format(root=ROOT, port=port))
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2"], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r}:\n{o[-3000:]}"


# ---------------------------------------------------------------------------
# the engines' own exchange (ptgpu_comm_init / ptgpu_shard_export + connect): kernels sum the
# per-tree scalars over peer-mapped mailboxes, no hook, no NCCL call
# ---------------------------------------------------------------------------
def _native_shards(tree, seqs, S, model, world, directional=True):
    from pt_b200.engine import Engine, comm_init
    from pt_b200.sharded import shard_range
    f, s, r = model
    n = tree.tip_count
    nbuf = 3 * (n - 2) if directional else n - 2
    engines = []
    for k in range(world):
        b, e_ = shard_range(S, world, k)
        e = Engine(n, nbuf, e_ - b, len(r), 2 * n - 3, nbuf)
        e.set_model(f, s, r)
        for tip in tree.nodes[:n]:
            e.set_tip_states(tip.clv_index, seqs[tip.label][b:e_])
        e.set_shard(b, S)
        engines.append(e)
    comm_init(engines)
    return engines


def _in_threads(engines, fn):
    outs, errs = [None] * len(engines), []

    def worker(k):
        try:
            outs[k] = fn(engines[k], k)
        except Exception as exc:  # pragma: no cover
            errs.append((k, exc))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(engines))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errs, errs
    return outs


@pytest.mark.parametrize("name,world", [("ds1", 2), ("ds1", 3), ("five", 2)])
def test_native_exchange_single_purpose_calls(name, world):
    """edge lnL, OptimizeBranch, pllmod's bounded Newton and the NNI batch on shards that
    exchange through peer memory == the unsharded engine"""
    from pt_b200.engine import ops_array
    from pt_b200.sharded import shard_range
    tree, seqs, g = load_set(name)
    model = model_of(g)
    n, S = tree.tip_count, g["sites"]
    if shard_range(S, world, world - 1)[0] >= S:
        pytest.skip("alignment has fewer tiles than shards")
    br = tree.branches()
    plan = ops_array(tree.directional_plan())
    ops, midx, mlen = tree.full_traversal()
    cands = cand_array(tree.nni_candidates())
    root = tree.by_index[g["root"]]
    rb = root.back

    def run(e, k=0):
        out = {}
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(ops))
        out["lnl"] = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index,
                                          rb.scaler_index, root.pmatrix_index)
        out["opt"] = e.optimize_branch(root.clv_index, root.scaler_index, rb.clv_index,
                                       rb.scaler_index, root.pmatrix_index, root.length)
        out["bounded"] = e.optimize_branch_bounded(root.clv_index, root.scaler_index, rb.clv_index,
                                                   rb.scaler_index, root.length)
        e.update_prob_matrices([root.pmatrix_index], [root.length])
        e.update_partials(plan)
        out["nni"] = e.score_nni_batch(cands)
        out["scalers"] = {i: e.get_scaler(i) for i in range(n - 2)}
        return out

    whole = make_engine(tree, seqs, *model, directional=True)
    ref = run(whole)
    whole.close()
    engines = _native_shards(tree, seqs, S, model, world)
    try:
        outs = _in_threads(engines, run)
        for o in outs:
            assert abs(o["lnl"] - ref["lnl"]) <= 1e-12 * abs(ref["lnl"])
            assert abs(o["lnl"] - fh(g["lnl"])) <= 1e-12 * abs(ref["lnl"])
            assert abs(o["opt"][0] - ref["opt"][0]) <= 1e-10 * max(1.0, ref["opt"][0])
            assert o["opt"][1] == ref["opt"][1]
            assert abs(o["bounded"][0] - ref["bounded"][0]) <= 1e-10 and o["bounded"][1:] == ref["bounded"][1:]
            assert np.allclose(o["nni"][0], ref["nni"][0], rtol=1e-9, atol=1e-12)
            assert np.allclose(o["nni"][1], ref["nni"][1], rtol=1e-12, atol=0)
            assert not o["nni"][2].any()
        for o in outs[1:]:   # every shard saw the same global sums, bit for bit
            assert o["lnl"] == outs[0]["lnl"] and o["opt"] == outs[0]["opt"]
            assert np.array_equal(o["nni"][1], outs[0]["nni"][1])
        for i in range(n - 2):
            assert np.array_equal(np.concatenate([o["scalers"][i] for o in outs]), ref["scalers"][i])
        status = [e.shard_status() for e in engines]
        assert all(st[0] == world and st[1] == 0 for st in status)
        assert len({st[2] for st in status}) == 1 and status[0][2] > 0   # same number of exchanges
    finally:
        for e in engines:
            e.close()


@pytest.mark.parametrize("name,world", [("ds1", 2), ("ds1", 3), ("five", 2)])
def test_native_exchange_whole_tree_program_on_shards(name, world):
    """Partition::OptimizeAllBranches as ONE persistent kernel per shard (the cross-shard sums of
    d1/d2 and lnL inside the kernel) == the unsharded kernel == the reference's golden result"""
    import all_branches as ab
    from pt_b200.engine import ops_array
    from pt_b200.sharded import shard_range
    from pytree import Tree
    tree, seqs, g = load_set(name)
    model = model_of(g)
    n, S = tree.tip_count, g["sites"]
    if shard_range(S, world, world - 1)[0] >= S:
        pytest.skip("alignment has fewer tiles than shards")
    newick = tree.newick()

    def run(e, k=0):
        t = Tree(newick)
        before = e.counters()["launches"]
        lnl0, lnl, passes = ab.visit(e, t)
        launches = e.counters()["launches"] - before
        return dict(lnl0=lnl0, lnl=lnl, passes=passes, lengths=t.branches(), launches=launches,
                    scalers={i: e.get_scaler(i) for i in range(n - 2)})

    whole = make_engine(tree, seqs, *model)
    ref = run(whole)
    whole.close()
    assert abs(ref["lnl"] - fh(g["oab_lnl"])) < 1e-9 * abs(ref["lnl"])
    engines = _native_shards(tree, seqs, S, model, world, directional=False)
    try:
        outs = _in_threads(engines, run)
        for o in outs:
            assert o["passes"] == ref["passes"]
            assert abs(o["lnl0"] - ref["lnl0"]) <= 1e-12 * abs(ref["lnl0"])
            assert abs(o["lnl"] - ref["lnl"]) <= 1e-12 * abs(ref["lnl"])
            assert abs(o["lnl"] - fh(g["oab_lnl"])) < 1e-9 * abs(o["lnl"])
            for pm, v in ref["lengths"].items():
                assert abs(o["lengths"][pm] - v) <= 1e-9 * max(1.0, v), pm
            for pm, v in g["oab_lengths"].items():
                assert abs(o["lengths"][int(pm)] - fh(v)) < 1e-6 * max(1.0, fh(v)), pm
        for o in outs[1:]:
            assert o["lnl"] == outs[0]["lnl"] and o["lengths"] == outs[0]["lengths"]
        for i in range(n - 2):
            assert np.array_equal(np.concatenate([o["scalers"][i] for o in outs]), ref["scalers"][i])
        # pmatrix + partials + edge lnL (kernel + exchange) + ONE step-program kernel
        assert outs[0]["launches"] <= 6, outs[0]["launches"]
    finally:
        for e in engines:
            e.close()


def test_native_exchange_times_out_instead_of_hanging(monkeypatch):
    """a shard whose peer never calls gets PTGPU_ERR_STATE after PTGPU_SHARD_TIMEOUT_MS"""
    from pt_b200.engine import PtGpuError, ops_array
    monkeypatch.setenv("PTGPU_SHARD_TIMEOUT_MS", "300")
    tree, seqs, g = load_set("five")
    model = model_of(g)
    engines = _native_shards(tree, seqs, g["sites"], model, 2, directional=False)
    try:
        e = engines[0]
        ops, midx, mlen = tree.full_traversal()
        e.update_prob_matrices(midx, mlen)
        e.update_partials(ops_array(ops))
        root = tree.by_index[g["root"]]
        rb = root.back
        with pytest.raises(PtGpuError) as err:
            e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                                 root.pmatrix_index)
        assert "did not answer" in str(err.value)
    finally:
        for e in engines:
            e.close()


PEER_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import torch, torch.distributed as dist
from common import fh, load_set, model_of
import all_branches as ab
from pt_b200.sharded import SiteShardedEngine
from pytree import Tree, cand_array
from pt_b200.engine import ops_array
rank = int(sys.argv[1]); world = int(sys.argv[2]); dev = int(sys.argv[3 + rank])
torch.cuda.set_device(dev)
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=world)
for name in {names!r}:
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    n = tree.tip_count
    e = SiteShardedEngine(n, 3 * (n - 2), g["sites"], len(r), 2 * n - 3, 3 * (n - 2), device=dev,
                          exchange="peer")
    e.set_model(f, s, r)
    for tip in tree.nodes[:n]:
        e.set_tip_states(tip.clv_index, seqs[tip.label])
    t = Tree(tree.newick())
    lnl0, lnl, passes = ab.visit(e, t)
    assert abs(lnl0 - fh(g["lnl"])) <= 1e-12 * abs(lnl0), (lnl0, fh(g["lnl"]))
    assert abs(lnl - fh(g["oab_lnl"])) < 1e-9 * abs(lnl), (lnl, fh(g["oab_lnl"]))
    for pm, v in g["oab_lengths"].items():
        assert abs(t.branches()[int(pm)] - fh(v)) < 1e-6 * max(1.0, fh(v)), pm
    if {full!r}:
        br = tree.branches()
        e.update_prob_matrices(list(br.keys()), list(br.values()))
        e.update_partials(ops_array(tree.directional_plan()))
        lens, lnls, st = e.score_nni_batch(cand_array(tree.nni_candidates()))
        for gl, ll in zip(g["nni"], lnls):
            assert abs(ll - fh(gl["lnl"])) <= 1e-10 * abs(ll), (ll, fh(gl["lnl"]))
        assert not st.any()
    nr, err, ex = e.engine.shard_status()
    assert nr == world and err == 0 and ex > 0
    e.close()
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
"""


def _run_peer_workers(tmp_path, devices, names, full):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "peer_worker.py"
    script.write_text(PEER_WORKER.format(root=ROOT, port=port, names=names, full=full))
    world = len(devices)
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(world)] + [str(d) for d in devices],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = [p.communicate(timeout=900)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r}:\n{o[-3000:]}"


def test_peer_exchange_two_processes_two_gpus(tmp_path):
    """one process per GPU, mailboxes mapped through CUDA IPC handles (what bench.py does
    under torchrun)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    _run_peer_workers(tmp_path, [0, 1], ("ds1", "five"), True)
