#!/bin/bash
# A/B of engine variants through environment switches: scripts/ab.sh "VAR=a" "VAR=b" ...
for cfg in "$@"; do
  echo "=== $cfg"
  env $cfg python bench.py --steps 5 --warmup 3 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): print(line[:200]); continue
    d=json.loads(line)
    print('ms_per_step %.2f value %.0f e2e %.0f' % (d['ms_per_step'], d['value'], d['e2e']['value']), {k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()}, 'chk', d['lnl_checksum'])
"
done
