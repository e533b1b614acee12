# ncu evidence for the step-program kernel (radius-1 batch), 64 taxa x 100k sites
set -x
CMD="python bench.py --radius 1 --sites 100000 --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/r1c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r1c_launches.csv $CMD > gpurun_out/r1c_ncu_launches.log 2>&1
echo launches rc=$?
$CMD > gpurun_out/r1c_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:blo_kernel' -s 1 -c 1 -o gpurun_out/r1c_prof -f $CMD > gpurun_out/r1c_ncu_full.log 2>&1
echo full rc=$?
tail -3 gpurun_out/r1c_ncu_full.log
tail -1 gpurun_out/r1c_plain2.log | cut -c1-300
ls -la gpurun_out | tail -8
