"""CPU tests of the multi-GPU plumbing (no GPU needed): shard ranges, and the
site-sharded driver over torch.distributed/gloo with world_size 2, using the
oracle-backed stand-in engine (tests/fake_engine.py).  The sharded result must
equal the reference's golden values for the unsharded alignment."""
import os
import socket
import subprocess
import sys

import pytest

from pt_b200.sharded import TILE_SITES, shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("sites,world", [(1949, 1), (1949, 2), (1949, 8), (1000000, 8), (255, 2),
                                          (100000, 4), (50000, 3)])
def test_shard_ranges_partition_the_alignment(sites, world):
    ranges = [shard_range(sites, world, r) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == sites
    for (b0, e0), (b1, e1) in zip(ranges, ranges[1:]):
        assert e0 == b1 and b0 <= e0
    for b, e in ranges:
        assert b % TILE_SITES == 0 or b == sites      # tile-aligned starts
    sizes = [e - b for b, e in ranges]
    assert max(sizes) - min(s for s in sizes if s) <= TILE_SITES * ((sites // TILE_SITES) // world + 1)


WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import torch.distributed as dist
from common import fh, fha, load_set, model_of
from fake_engine import FakeEngine
from pt_b200.engine import ops_array
from pt_b200.sharded import SiteShardedEngine, shard_range

dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size=2)
for name in ("ds1", "five"):
    tree, seqs, g = load_set(name)
    f, s, r = model_of(g)
    n = tree.tip_count
    e = SiteShardedEngine(n, n - 2, g["sites"], len(r), 2 * n - 3, n - 2, engine_factory=FakeEngine)
    assert (e.begin, e.end) == shard_range(g["sites"], 2, dist.get_rank()) and e.end > e.begin
    e.set_model(f, s, r)
    for tip in tree.nodes[:n]:
        e.set_tip_states(tip.clv_index, seqs[tip.label])
    ops, midx, mlen = tree.full_traversal()
    e.update_prob_matrices(midx, mlen)
    e.update_partials(ops)
    root = tree.by_index[g["root"]]; rb = root.back
    lnl = e.edge_loglikelihood(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                               root.pmatrix_index)
    assert abs(lnl - fh(g["lnl"])) <= 1e-12 * abs(lnl), (lnl, fh(g["lnl"]))
    # per-site scaler counts are exactly libpll's, shard by shard
    for i, nz in g["scalers_nonzero"].items():
        exp = np.zeros(g["sites"], dtype=np.uint32)
        for site, cnt in nz:
            exp[site] = cnt
        assert np.array_equal(e.gather_scaler(int(i)), exp)
    ob = g["optimize_branch"][0]
    before = e.allreduce_calls
    ln, iters = e.optimize_branch(root.clv_index, root.scaler_index, rb.clv_index, rb.scaler_index,
                                  root.pmatrix_index, fh(ob["len0"]))
    assert abs(ln - fh(ob["len"])) <= 1e-9 * max(1.0, ln) and iters == ob["iters"], (ln, iters, ob)
    # one [2]-double allreduce per Newton step, nothing else crosses ranks
    assert e.allreduce_calls - before == iters
    e.close()
dist.barrier()
dist.destroy_process_group()
print("rank", sys.argv[1], "ok")
"""


def test_site_sharded_driver_world2_gloo(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, port=port))
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r}:\n{o[-3000:]}"
        assert f"rank {r} ok" in o
