# round 2: launch list + one --set full capture of the three hot kernels of the headline step
# (run under gpurun; each ncu pass only after the same command exited 0 without ncu)
set -x
TAG=${1:-r2}
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity-at-size --no-sharded"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo launches rc=$?
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:partials_kernel_v4|nni_prepare_kernel_v2|newton_kernel' -s 3 -c 3 -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo full rc=$?
tail -3 gpurun_out/${TAG}_ncu_full.log
