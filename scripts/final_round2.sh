# round 2, final 1-GPU evidence: the whole GPU suite, then the bench lines (own arm, reference arm,
# the named shapes, the DS1 walk) and the walk timings.  Run under gpurun.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r2_final_tests.log 2>&1; echo tests rc=$?; tail -4 gpurun_out/r2_final_tests.log
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; echo ref rc=$?
timeout 900 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?
timeout 600 python bench.py --workload cand100k --steps 2 --warmup 1 > gpurun_out/r2_cand100k_n1.json 2> gpurun_out/r2_cand100k_n1.err; echo cand rc=$?
timeout 600 python bench.py --workload stress1000 --steps 3 --warmup 2 > gpurun_out/r2_stress1000_n1.json 2> gpurun_out/r2_stress1000_n1.err; echo stress rc=$?
timeout 300 python bench.py --workload ptw-ds1 --wanderers 1 --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ptw_ds1_w1.json 2> gpurun_out/r2_ptw_w1.err; echo ptw1 rc=$?
timeout 300 python bench.py --workload ptw-ds1 --wanderers 16 --steps 3 --warmup 1 > gpurun_out/r2_ptw_ds1_w16.json 2> gpurun_out/r2_ptw_w16.err; echo ptw16 rc=$?
timeout 600 python scripts/walk_timing.py quick > gpurun_out/r2_walk_timing.log 2>&1; tail -20 gpurun_out/r2_walk_timing.log
for f in gpurun_out/r2_bench_n1.json gpurun_out/r2_bench_reference.json gpurun_out/r2_cand100k_n1.json gpurun_out/r2_stress1000_n1.json gpurun_out/r2_ptw_ds1_w1.json gpurun_out/r2_ptw_ds1_w16.json; do echo $f; cut -c1-330 $f | tail -1; done
