# round 2, one 8-GPU box: the named shapes and modes at N = 8, 4, 2 (N = 1 lines come from 1-GPU calls)
run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N "$@"; }
timeout 300 python -m pytest tests/test_gpu_sharded.py -m gpu -q -k "two_gpus or two_processes" > gpurun_out/r2_tests_2gpu.log 2>&1; tail -2 gpurun_out/r2_tests_2gpu.log
PTGPU_SITE_SHARDS=8 timeout 300 python -m pytest tests/test_gpu_dropin.py -m gpu -q -k "ptpll_on_site" > gpurun_out/r2_tests_shards8.log 2>&1; tail -2 gpurun_out/r2_tests_shards8.log
for N in 8 4 2; do
  timeout 400 bash -c "$(declare -f run); run $N --steps 5 --warmup 3" > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
  timeout 400 bash -c "$(declare -f run); run $N --workload cand100k --steps 2 --warmup 1" > gpurun_out/r2_cand100k_n$N.json 2> gpurun_out/r2_cand100k_n$N.err
  timeout 300 bash -c "$(declare -f run); run $N --workload stress1000 --steps 3 --warmup 2" > gpurun_out/r2_stress1000_n$N.json 2> gpurun_out/r2_stress1000_n$N.err
  timeout 300 python bench.py --workload ptw-ds1 --wanderers 16 --gpus $N --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ptw_ds1_w16_n$N.json 2> gpurun_out/r2_ptw_n$N.err
  tail -q -n1 gpurun_out/r2_*_n$N.err | cut -c1-200
done
ls -la gpurun_out | grep "_n[248].json"
