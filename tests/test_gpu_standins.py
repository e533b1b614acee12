"""The topography walk on stand-ins for the hohna data sets that are not in the reference
repository (published shapes DS3 36 x 1812, DS6 50 x 1133, DS8 64 x 1008; alignments simulated
under DS1's model, tests/golden/make_standin_golden.py): ptw_batched on one B200 must produce
the good / visited / tested tables the UNMODIFIED reference's ptw_threads produced on the
same files (tests/golden/standin_golden.json) -- the walk's parity at other tree sizes than
DS1's 27 taxa, up to the 64 taxa of the headline workload.  Radius 0 goes through
ptgpu_score_nni_batch, radius 1 through ptgpu_optimize_local_batch; every visit through
ptgpu_optimize_all_branches."""
import json
import math
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
    tables = run(exe, name, gold["args"])
    got = summarise(tables)
    assert got["counts"] == gold["counts"]
    for key in ("good_topologies_sha", "visited_topologies_sha", "tested_topologies_sha"):
        assert got[key] == gold[key], key
    # scores are printed with 6 significant digits (src/ptw_threads.cpp:28)
    from make_ptw_golden import digest, score_buckets
    if "good" in gold:
        assert got["good"] == gold["good"]
    else:   # thousands of good trees: the golden file keeps their digest and per-bucket score sums
        got["good_sha"], got["good_buckets"] = digest(got["good"]), score_buckets(got["good"])
    for table in ("good", "visited", "tested"):
        if table + "_sha" not in gold or got[table + "_sha"] == gold[table + "_sha"]:
            continue
        # Same trees (the topology digests above), another digest of the lines: with several 10^5 scores
        # printed to 6 significant digits a few sit on a rounding edge, where a 1e-12 relative difference
        # in lnL flips the last printed digit.  Goldens of the large sets carry per-bucket score sums
        # (make_ptw_golden.score_buckets): at most a handful of buckets may differ, by one last-digit unit.
        assert table + "_buckets" in gold, f"{table} scores differ"
        unit = 10.0 ** (math.floor(math.log10(gold["score_scale"])) - 5)
        off = [abs(a - b) for a, b in zip(got[table + "_buckets"], gold[table + "_buckets"]) if abs(a - b) > 1e-6]
        assert len(off) <= 5 and max(off) <= 1.5 * unit, (table, len(off), max(off), unit)
