#!/bin/bash
# A/B of two engine builds on the same box: PTGPU_LIB selects the library bench.py loads.
mkdir -p gpurun_out
show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(sys.argv[2], round(d['value'],1), round(d['ms_per_step'],2), {k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()}, 'e2e', round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['step_host_ms'], d['e2e_step_host_ms'])" $1 "$2"; }
for rep in 1 2 3; do
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab_new.json 2>gpurun_out/ab_new.err; show gpurun_out/ab_new.json new
  PTGPU_BENCH_NO_SAMPLER=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab_new1.json 2>gpurun_out/ab_new1.err; show gpurun_out/ab_new1.json no_sampler
done
