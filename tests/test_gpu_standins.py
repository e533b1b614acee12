"""The topography walk on stand-ins for the hohna data sets that are not in the reference
repository (published shapes DS3 36 x 1812, DS6 50 x 1133, DS8 64 x 1008; alignments simulated
under DS1's model, tests/golden/make_standin_golden.py): ptw_batched on one B200 must produce
the good / visited / tested tables the UNMODIFIED reference's ptw_threads produced on the
same files (tests/golden/standin_golden.json) -- the walk's parity at other tree sizes than
DS1's 27 taxa, up to the 64 taxa of the headline workload.  Radius 0 goes through
ptgpu_score_nni_batch, radius 1 through ptgpu_optimize_local_batch; every visit through
ptgpu_optimize_all_branches."""
import json
import os
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))

with open(os.path.join(HERE, "golden", "standin_golden.json")) as _f:
    GOLD = json.load(_f)


@pytest.mark.parametrize("name", sorted(GOLD))
def test_walk_tables_equal_the_reference_s(name):
    from make_ptw_golden import summarise
    from make_standin_golden import run
    exe = os.path.join(ROOT, "pt_b200", "bin", "ptw_batched")
    if not os.path.exists(exe):
        pytest.fail("pt_b200/bin/ptw_batched not built (pt_b200/host/Makefile drivers needs a pt checkout at build time)")
    gold = GOLD[name]
    got = summarise(run(exe, name, gold["args"]))
    assert got["counts"] == gold["counts"]
    for key in ("good_topologies_sha", "visited_topologies_sha", "tested_topologies_sha"):
        assert got[key] == gold[key], key
    # scores are printed with 6 significant digits (src/ptw_threads.cpp:28)
    assert got["good"] == gold["good"]
    assert got["visited_sha"] == gold["visited_sha"]
    assert got["tested_sha"] == gold["tested_sha"]
