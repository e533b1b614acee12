#!/bin/bash
# BASELINE configs[2]: site-sharded full-tree lnL + OptimizeAllBranches at N = 1, 2, ... GPUs of one box
mkdir -p gpurun_out
for N in "$@"; do
  if [ "$N" = 1 ]; then
    python bench.py --workload sharded --gpus 1 --steps 3 --warmup 3 > gpurun_out/sharded_n1.json 2> gpurun_out/sharded_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --workload sharded --gpus $N --steps 3 --warmup 3 > gpurun_out/sharded_n$N.json 2> gpurun_out/sharded_n$N.err
  fi
  echo "N=$N rc=$?"; tail -2 gpurun_out/sharded_n$N.err
  python -c "
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(d['n_gpus'], round(d['value'],3), 'trees/s', round(d['ms_per_step'],1), 'ms', 'e2e', round(d['e2e']['value'],3), 'passes', d['optimize_all_branches_passes'], repr(d['lnl_optimised']), {k:(round(v['ms_total_per_step'],1), round(v.get('achieved_gbs',0))) for k,v in d['kernels'].items()}, 'gap', round(d['outside_kernels_share'],2), 'allreduce/step', d['allreduce_calls_per_step'])
" gpurun_out/sharded_n$N.json
done
