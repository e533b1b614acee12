#!/bin/bash
# one ncu --set full capture of the single-branch Newton sweep at 1M sites: 6-stage ring vs direct loads
mkdir -p gpurun_out
CMD="python bench.py --workload sharded --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/ring_plain.log 2>&1 || exit 1
for R in 1 0; do
  PTGPU_FEW_WORK_RING=$R timeout 600 ncu --set full --clock-control none -k regex:newton_sweep_kernel -s 200 -c 3 \
     -o gpurun_out/r1e_ring$R -f $CMD > gpurun_out/ring_ncu$R.log 2>&1
  echo "ring=$R rc=$?"
done
