set -x
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/r1e_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1e_launches.csv $CMD > gpurun_out/r1e_ncu_launches.log 2>&1
echo launches rc=$?
$CMD > gpurun_out/r1e_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:partials_kernel_v4|nni_prepare_kernel_v2|newton_kernel' -s 3 -c 3 -o gpurun_out/r1e_prof -f $CMD > gpurun_out/r1e_ncu_full.log 2>&1
echo full rc=$?
tail -3 gpurun_out/r1e_ncu_full.log
ls -la gpurun_out
