# round 2, last 1-GPU evidence: the whole GPU suite, then the headline bench line (with the re-stamped traffic)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2_final_tests.log 2>&1; echo tests rc=$?; tail -14 gpurun_out/r2_final_tests.log
timeout 600 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?
tail -1 gpurun_out/r2_bench_n1.json | cut -c1-300
timeout 200 python bench.py --workload ptw-ds1 --wanderers 1 --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ptw_ds1_w1.json 2> gpurun_out/r2_ptw_w1.err; echo ptw1 rc=$?
tail -1 gpurun_out/r2_ptw_ds1_w1.json | cut -c1-200
