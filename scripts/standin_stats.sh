#!/bin/bash
D=tests/golden/data/standins/DS8
A="--raxml-info $D/RAxML_info.DS8 --lnl-offset -1.5 --optimization-radius 0"
for T in 1 16; do
  s=$(date +%s%N)
  PTGPU_HOST_STATS=1 oracle/_ref/gpu/ptw_batched $A --threads $T --polling-time 1 $D/DS8.tre $D/DS8.fasta gpurun_out/walk_good.tsv 2>&1 | sort | uniq -c | sort -rn | head -8
  e=$(date +%s%N); echo "T=$T wall_ms=$(( (e - s) / 1000000 ))"
done
