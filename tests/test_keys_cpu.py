"""Topology keys: this repository's implementation of pt::CompressedTree
(pt_b200/host/ptw/compressed_tree.cpp: integer ranks, no clone, no string compares) against the
reference's own src/compressed_tree.cpp -- same bits, same label indices, same hash, same decoded
Newick on random trees and on chains of NNI moves, at 5 to 200 taxa.  CPU only."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "keys_ref")
OURS = os.path.join(ROOT, "oracle", "_ref", "keys_ours")


def trees_file(tmp_path, taxa, count, seed):
    from pt_b200.pytree import Tree, random_tree_newick
    rng = np.random.default_rng(seed)
    lines = []
    t = Tree(random_tree_newick(taxa, rng))
    for k in range(count):
        if k % 7 == 0:   # a fresh topology now and then (same labels), NNI moves in between
            t = Tree(random_tree_newick(taxa, rng))
        else:
            inner = [n for n in t.nodes[t.tip_count:] if not n.back.is_tip]
            if inner:
                t.nni(inner[int(rng.integers(len(inner)))], int(rng.integers(1, 3)))
        lines.append(t.newick())
    p = tmp_path / f"trees_{taxa}.nwk"
    p.write_text("\n".join(lines) + "\n")
    return str(p)


def need_binaries():
    if not (os.path.exists(REF) and os.path.exists(OURS)):
        pytest.skip("oracle/_ref/keys_{ref,ours} not built (oracle/Makefile.keys needs /root/reference)")


@pytest.mark.parametrize("taxa,count", [(5, 60), (27, 120), (64, 120), (200, 40)])
def test_keys_equal_the_reference_s(tmp_path, taxa, count):
    need_binaries()
    path = trees_file(tmp_path, taxa, count, 1000 + taxa)
    a = subprocess.run([REF, path], capture_output=True, text=True, check=True).stdout
    b = subprocess.run([OURS, path], capture_output=True, text=True, check=True).stdout
    assert a == b
    assert a.count(";") == 3 * count   # every tree twice in a row, then once more


def test_key_construction_is_faster_than_the_reference_s(tmp_path):
    need_binaries()
    path = trees_file(tmp_path, 64, 50, 7)
    t = {}
    for name, exe in (("ref", REF), ("ours", OURS)):
        r = subprocess.run([exe, path, "200"], capture_output=True, text=True, check=True)
        t[name] = float(r.stderr.split()[0])
    # the authority's pattern: the key of one tree two or three times in a row
    r = subprocess.run([OURS, path, "200", "2"], capture_output=True, text=True, check=True)
    t["ours_twice"] = float(r.stderr.split()[0])
    print(f"us per key at 64 taxa: reference {t['ref']:.2f}, this repository {t['ours']:.2f} "
          f"({t['ours_twice']:.2f} when each tree is asked for twice running)")
    assert t["ours"] < t["ref"]
    assert t["ours_twice"] < t["ours"]
