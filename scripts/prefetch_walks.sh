#!/bin/bash
# DS1 (config 2) and the DS8 stand-in walk with visits computed ahead of time (PTGPU_PREFETCH_VISITS=n)
mkdir -p gpurun_out
run() { # name threads prefetch args...
  local name=$1 T=$2 P=$3; shift 3
  local s=$(date +%s%N)
  PTGPU_PREFETCH_VISITS=$P oracle/_ref/gpu/ptw_batched "$@" --threads $T --polling-time 1 --tested-trees gpurun_out/pf_tested.tsv gpurun_out/pf_good.tsv 2> gpurun_out/pf.err
  local rc=$? e=$(date +%s%N)
  echo "$name threads=$T prefetch=$P rc=$rc wall_ms=$(( (e - s) / 1000000 )) good=$(wc -l < gpurun_out/pf_good.tsv) tested=$(wc -l < gpurun_out/pf_tested.tsv) sha=$(sort gpurun_out/pf_good.tsv | sha256sum | cut -c1-12) $(grep -h 'visits ahead' gpurun_out/pf.err)"
}
D1=tests/golden/data/ds1; D8=tests/golden/data/standins/DS8
A1="--raxml-info $D1/RAxML_info.DS1 --lnl-offset -5 --optimization-radius 1 $D1/RAxML_bestTree.DS1 $D1/DS1.fasta"
A8="--raxml-info $D8/RAxML_info.DS8 --lnl-offset -1.5 --optimization-radius 0 $D8/DS8.tre $D8/DS8.fasta"
if [ "$1" = multi ]; then
  for P in 0 4 0 4; do run ds1 16 $P $A1; done
  for P in 0 4 8 0 4 8; do run ds8 16 $P $A8; done
  exit 0
fi
for P in 0 2 4 8; do run ds1 1 $P $A1; done
for P in 0 4 8 16; do run ds8 1 $P $A8; done
run ds1 8 4 $A1; run ds8 8 8 $A8; run ds8 16 8 $A8
