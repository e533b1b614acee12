# round 2, final: the three named shapes at N = 8 on one 8-GPU box with the round's last kernels
run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N "$@"; }
N=8
timeout 400 bash -c "$(declare -f run); run $N --steps 5 --warmup 3" > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
timeout 300 bash -c "$(declare -f run); run $N --workload cand100k --steps 2 --warmup 1" > gpurun_out/r2_cand100k_n$N.json 2> gpurun_out/r2_cand100k_n$N.err
timeout 300 bash -c "$(declare -f run); run $N --workload stress1000 --steps 3 --warmup 2" > gpurun_out/r2_stress1000_n$N.json 2> gpurun_out/r2_stress1000_n$N.err
tail -q -n1 gpurun_out/r2_*_n$N.err | cut -c1-200
for f in gpurun_out/r2_bench_n$N.json gpurun_out/r2_cand100k_n$N.json gpurun_out/r2_stress1000_n$N.json; do tail -1 $f | cut -c1-250; done
