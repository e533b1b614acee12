"""DS1 config-2 and DS8-shape walks, one wanderer, for a few speculation depths / worker counts (run under gpurun)."""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from walk_timing import DS1, DS8, B, run   # noqa: E402 (runs its own table first unless told not to)
for rep in range(2):
    for depth in (0, 1, 2, 3, 6):
        run(f"ds1 depth {depth}", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_DEPTH=depth)
    run("ds1 depth 3 workers 10", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_DEPTH=3, PTGPU_PREFETCH_VISITS=10)
for depth in (0, 1, 3):
    run(f"ds8 depth {depth}", B + "ptw_batched", 1, DS8, PTGPU_PREFETCH_DEPTH=depth)
