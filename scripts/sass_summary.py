"""SASS opcode summary per kernel of pt_b200/lib/libptgpu.so (cuobjdump -sass; no GPU needed):
counts of the Blackwell-specific mnemonics the kernels rely on and of the FP64 arithmetic.
    python scripts/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pt_b200", "lib", "libptgpu.so")
WATCH = ["UBLKCP", "SYNCS", "LDGSTS", "LDG.E.ENL2.256", "LDG.E.256", "STG.E.ENL2.256", "STG.E.256", "LDS.128",
         "UCGABAR", "BAR.SYNC", "SHFL", "DFMA", "DMUL", "DADD", "MUFU", "ATOM", "RED", "UTC", "HMMA", "LDL", "STL"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
kern, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        counts[kern]["instructions"] += 1
        for w in WATCH:
            if op.startswith(w) or (w in ("LDG.E.ENL2.256", "STG.E.ENL2.256", "LDG.E.256", "STG.E.256", "LDS.128") and w in op):
                counts[kern][w] += 1
print(f"{os.path.relpath(LIB, ROOT)}: cubin architectures {arch}")
print("No UTC*MMA / HMMA is expected: the per-site 4x4 products are not a dense contraction (north_star).")
print("UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier, LDGSTS = cp.async, UCGABAR = cluster barrier.\n")
hot = ("partials_kernel_v4<4>", "nni_prepare_kernel_v2<4>", "newton_kernel<4, true, 0>", "newton_kernel<4, true, 6>",
       "blo_kernel<4, false>", "blo_kernel<4, true>", "edge_lnl_kernel<4>", "sumtable_kernel<4>", "pmatrix_kernel<4>",
       "shard_allreduce_kernel")
for k, c in counts.items():
    name = k.replace("void ", "").replace("ptgpu::", "")
    if not any(name.startswith(h) or name == h for h in hot):
        continue
    cols = "  ".join(f"{w}={c[w]}" for w in WATCH if c[w])
    print(f"{name:34s} {c['instructions']:6d} instr  {cols}")
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print("\nwhole library:", "  ".join(f"{w}={tot[w]}" for w in WATCH))
