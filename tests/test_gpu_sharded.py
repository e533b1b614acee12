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
