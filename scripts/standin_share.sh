#!/bin/bash
D=tests/golden/data/standins/DS8
A="--raxml-info $D/RAxML_info.DS8 --lnl-offset -1.5 --optimization-radius 0"
for M in 1 2 4 8 16; do for rep in 1 2; do
  s=$(date +%s%N)
  PTGPU_COOP_SHARE_MAX=$M oracle/_ref/gpu/ptw_batched $A --threads 16 --polling-time 1 $D/DS8.tre $D/DS8.fasta gpurun_out/walk_good.tsv 2>/dev/null
  e=$(date +%s%N); echo "share_max=$M wall_ms=$(( (e - s) / 1000000 )) good=$(wc -l < gpurun_out/walk_good.tsv)"
done; done
