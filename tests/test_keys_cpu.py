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


# ---- the 16-bit label index (pt_b200/host/ptw/wide_label_dictionary.hpp) ----------------------
WIDE = os.path.join(ROOT, "oracle", "_ref", "keys_wide")


def splits_of(newick):
    """the non-trivial bipartitions of a Newick tree, each as the frozenset of tip labels on the
    side that does not hold the alphabetically smallest label"""
    from pt_b200.pytree import Tree
    t = Tree(newick)
    labels = sorted(t.labels())
    anchor = labels[0]
    out = set()

    def below(n):   # tip labels behind record n (iterative post-order: 1000-taxon trees are deep)
        stack, seen = [(n, False)], {}
        while stack:
            r, done = stack.pop()
            if r.is_tip:
                seen[id(r)] = frozenset([r.label])
            elif done:
                seen[id(r)] = seen[id(r.next.back)] | seen[id(r.next.next.back)]
                side = seen[id(r)]
                side = side if anchor not in side else frozenset(labels) - side
                if 1 < len(side) < len(labels) - 1:
                    out.add(side)
            else:
                stack += [(r, True), (r.next.back, False), (r.next.next.back, False)]
        return seen[id(n)]

    r = t.root
    for x in (r, r.next, r.next.next):
        below(x.back)
    return out


def rotated_newick(tree, rng):
    """the same unrooted tree written from another inner node, children in random order"""
    inner = [n for n in tree.nodes[tree.tip_count:]]
    root = inner[int(rng.integers(len(inner)))]

    def rec(n):
        stack, text = [(n, 0)], {}
        while stack:
            r, state = stack.pop()
            if r.is_tip:
                text[id(r)] = f"{r.label}:{r.length!r}"
            elif state == 0:
                stack += [(r, 1), (r.next.back, 0), (r.next.next.back, 0)]
            else:
                kids = [text[id(r.next.back)], text[id(r.next.next.back)]]
                if rng.integers(2):
                    kids.reverse()
                text[id(r)] = "(" + ",".join(kids) + f"):{r.length!r}"
        return text[id(n)]

    parts = [rec(x.back) for x in (root, root.next, root.next.next)]
    rng.shuffle(parts)
    return "(" + ",".join(parts) + ");"


def key_blocks(stdout, count):
    """keys_dump prints 4 lines per key (bits, label indices, hash, decoded Newick), every tree
    twice in a row and then all once more: the first key of each tree"""
    lines = stdout.splitlines()
    assert len(lines) == 4 * 3 * count
    return [tuple(lines[8 * i: 8 * i + 4]) for i in range(count)]


@pytest.mark.parametrize("taxa,count", [(27, 60), (200, 40), (256, 20)])
def test_wide_label_index_changes_nothing_up_to_256_taxa(tmp_path, taxa, count):
    """built with the 16-bit dictionary, the keys of trees the reference can handle are the
    reference's: bits, indices, hash and decoded Newick"""
    need_binaries()
    if not os.path.exists(WIDE):
        pytest.skip("oracle/_ref/keys_wide not built")
    path = trees_file(tmp_path, taxa, count, 4000 + taxa)
    a = subprocess.run([REF, path], capture_output=True, text=True, check=True).stdout
    b = subprocess.run([WIDE, path], capture_output=True, text=True, check=True).stdout
    assert a == b


@pytest.mark.parametrize("taxa", [257, 300, 1000])
def test_wide_label_index_beyond_256_taxa(tmp_path, taxa):
    """more taxa than pt's uint8_t index allows (the reference throws std::length_error: no
    reference output exists): the key is a function of the unrooted topology alone -- equal for
    the same tree written from another node with its children in another order, different for
    NNI neighbours -- and decodes to that topology"""
    from pt_b200.pytree import Tree
    if not os.path.exists(WIDE):
        pytest.skip("oracle/_ref/keys_wide not built")
    count = 24
    path = trees_file(tmp_path, taxa, count, 9000 + taxa)
    rng = np.random.default_rng(taxa)
    originals = open(path).read().splitlines()
    # the first tree stays first (its tip order defines the dictionary); rotations appended
    rotated = [rotated_newick(Tree(nw), rng) for nw in originals]
    both = tmp_path / "both.nwk"
    both.write_text("\n".join(originals + rotated) + "\n")
    r = subprocess.run([WIDE, str(both)], capture_output=True, text=True, check=True)
    keys = key_blocks(r.stdout, 2 * count)
    for i in range(count):
        assert keys[i] == keys[count + i], f"tree {i}: another rotation, another key"
        assert splits_of(keys[i][3]) == splits_of(originals[i]), f"tree {i}: decoded topology differs"
    distinct = {frozenset(splits_of(nw)) for nw in originals}
    assert len({k[:2] for k in keys[:count]}) == len(distinct)
    assert max(int(x) for x in keys[0][1].strip("[]").split(",")) == taxa - 1   # indices beyond 255 in use
    if os.path.exists(REF):
        ref = subprocess.run([REF, path], capture_output=True, text=True)
        assert ref.returncode != 0   # "index type is not large enough"
