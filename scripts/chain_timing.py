"""One-wanderer walks with the sibling chains (run under gpurun)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from walk_timing import DS1, DS8, B, run   # noqa: E402
for rep in range(2):
    run("ds1 default (1 wanderer)", B + "ptw_batched", 1, DS1)
    run("ds8 default (1 wanderer)", B + "ptw_batched", 1, DS8)
run("ds8 no prefetch", B + "ptw_batched", 1, DS8, PTGPU_PREFETCH_VISITS=0)
run("ds8 workers 10", B + "ptw_batched", 1, DS8, PTGPU_PREFETCH_VISITS=10)
run("ds8 16 wanderers", B + "ptw_batched", 16, DS8)
