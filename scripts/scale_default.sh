#!/bin/bash
# the default (candidate-parallel) bench at several N on one box
mkdir -p gpurun_out
for N in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/scale_n$N.err | grep '^{' > gpurun_out/scale_n$N.json
  python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), d['clocks'])" gpurun_out/scale_n$N.json
done
