#!/bin/bash
# bench the engine variants built as pt_b200/lib/libptgpu_<name>.so (kernel tuning experiments)
for lib in "" $(ls pt_b200/lib/libptgpu_*.so 2>/dev/null); do
  echo "=== ${lib:-default}"
  PTGPU_LIB=$([ -n "$lib" ] && echo $PWD/$lib) python bench.py --steps 3 --warmup 3 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | python -c "
import sys, json
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): print(line[:200]); continue
    d=json.loads(line)
    print('ms_per_step %.2f value %.0f e2e %.0f' % (d['ms_per_step'], d['value'], d['e2e']['value']), {k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()}, 'chk', d['lnl_checksum'])
"
done
