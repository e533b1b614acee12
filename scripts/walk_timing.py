"""Wall time of the DS1 config-2 walk and the DS8-shape stand-in under variants of the host-side
machinery (run under gpurun): `label seconds | prefetch summary` lines."""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
D1 = "tests/golden/data/ds1"
S8 = "tests/golden/data/standins/DS8"
DS1 = ["--raxml-info", f"{D1}/RAxML_info.DS1", "--lnl-offset", "-5", "--optimization-radius", "1",
       "--tested-trees", "/tmp/t.tsv", f"{D1}/RAxML_bestTree.DS1", f"{D1}/DS1.fasta", "/tmp/g.tsv"]
DS8 = ["--raxml-info", f"{S8}/RAxML_info.DS8", "--lnl-offset", "-1.5", "--optimization-radius", "0",
       "--tested-trees", "/tmp/t.tsv", f"{S8}/DS8.tre", f"{S8}/DS8.fasta", "/tmp/g.tsv"]
B = "pt_b200/bin/"


def run(label, binary, threads, data, **env):
    e = dict(os.environ, **{k: str(v) for k, v in env.items()})
    cmd = [binary, "--threads", str(threads)] + (["--polling-time", "1"] if threads > 1 else []) + data
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, env=e)
    dt = time.perf_counter() - t0
    note = [l for l in r.stderr.splitlines() if l.startswith("visits")]
    print(f"{label:32s} {dt:7.2f} s  rc={r.returncode} | {note[0][22:260] if note else ''}", flush=True)


def main():
  quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
  for rep in range(1 if quick else 2):
      run("ds1 default (1 wanderer)", B + "ptw_batched", 1, DS1)
      run("ds1 reference keys", B + "ptw_batched_refkeys", 1, DS1)
      run("ds1 no prefetch", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_VISITS=0)
      run("ds1 depth 0", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_DEPTH=0)
      run("ds1 depth 2", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_DEPTH=2)
      run("ds1 workers 3", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_VISITS=3)
      run("ds1 workers 12", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_VISITS=12)
      run("ds1 no neighbourhoods", B + "ptw_batched", 1, DS1, PTGPU_PREFETCH_NEIGHBORHOODS=0)
  run("ds1 4 wanderers", B + "ptw_batched", 4, DS1)
  run("ds1 16 wanderers", B + "ptw_batched", 16, DS1)
  run("ds1 reference, 16 threads", "oracle/_ref/ptw_threads", 16, DS1)
  run("ds8 default (1 wanderer)", B + "ptw_batched", 1, DS8)
  run("ds8 reference keys", B + "ptw_batched_refkeys", 1, DS8)
  run("ds8 no prefetch", B + "ptw_batched", 1, DS8, PTGPU_PREFETCH_VISITS=0)
  run("ds8 depth 0", B + "ptw_batched", 1, DS8, PTGPU_PREFETCH_DEPTH=0)
  run("ds8 16 wanderers", B + "ptw_batched", 16, DS8)
  run("ds8 reference, 16 threads", "oracle/_ref/ptw_threads", 16, DS8)



if __name__ == "__main__":
    main()
