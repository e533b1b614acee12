#!/usr/bin/env python
"""Golden good_trees.tsv for BASELINE configs 1 and 2, produced by the UNMODIFIED
reference (oracle/_ref/ptw_threads, built by oracle/Makefile.ref) at --threads 1.

Run here (needs /root/reference at build time); the GPU tests compare
pt_b200/bin/ptw_threads_gpu (same pt sources, GPU-backed Partition) with the
file this writes: tests/golden/ptw_golden.json.  Lines are compared SORTED:
their order in the output is cuckoo-table iteration order.
"""
import hashlib
import json
import os
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
DS1 = os.path.join(ROOT, "tests", "golden", "data", "ds1")
CONFIGS = {
    # BASELINE.json configs[0]
    "config1": ["--lnl-offset", "-2", "--optimization-radius", "0"],
    # BASELINE.json configs[1] on DS1 (DS2-DS8 are not in the reference repo)
    "config2": ["--lnl-offset", "-5", "--optimization-radius", "1"],
}


def run(binary, extra, threads=1):
    with tempfile.TemporaryDirectory() as d:
        out = {k: os.path.join(d, k + ".tsv") for k in ("good", "visited", "tested")}
        cmd = [binary, "--raxml-info", os.path.join(DS1, "RAxML_info.DS1"), "--threads", str(threads),
               "--visited-trees", out["visited"], "--tested-trees", out["tested"]] + extra + [
               os.path.join(DS1, "RAxML_bestTree.DS1"), os.path.join(DS1, "DS1.fasta"), out["good"]]
        subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
        return {k: sorted(open(p).read().splitlines()) for k, p in out.items()}


def digest(lines):
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()


def score_buckets(lines, n=256):
    """per-bucket sums of the printed scores, trees dealt to n buckets by a hash of their Newick string:
    a table of several 10^5 scores printed with 6 significant digits is expected to hold a few that sit
    on a rounding edge (a 1e-12 relative difference in lnL then flips the last printed digit); such a
    flip moves ONE bucket by ONE unit of the last digit, anything systematic moves them all"""
    sums = [0.0] * n
    for l in lines:
        score, newick = l.split("\t", 1)
        sums[int.from_bytes(hashlib.sha1(newick.encode()).digest()[:4], "little") % n] += float(score)
    return [round(x, 6) for x in sums]


def summarise(tables):
    return dict(good=tables["good"], counts={k: len(v) for k, v in tables.items()},
                visited_buckets=score_buckets(tables["visited"]), tested_buckets=score_buckets(tables["tested"]),
                visited_sha=digest(tables["visited"]), tested_sha=digest(tables["tested"]),
                # topologies only (lnl column dropped): what "identical good_trees" pins even
                # if a printed 6-significant-digit lnl sat on a rounding edge
                good_topologies_sha=digest(sorted(l.split("\t", 1)[1] for l in tables["good"])),
                visited_topologies_sha=digest(sorted(l.split("\t", 1)[1] for l in tables["visited"])),
                tested_topologies_sha=digest(sorted(l.split("\t", 1)[1] for l in tables["tested"])))


if __name__ == "__main__":
    ref = os.path.join(ROOT, "oracle", "_ref", "ptw_threads")
    g = {name: dict(args=extra, **summarise(run(ref, extra))) for name, extra in CONFIGS.items()}
    with open(os.path.join(ROOT, "tests", "golden", "ptw_golden.json"), "w") as f:
        json.dump(g, f, indent=0)
    for k, v in g.items():
        print(k, v["counts"])
