#!/bin/bash
# Where does the DS1 walk spend its wall time?  Fixed cost (offset 0 = start tree only) vs the walk.
mkdir -p gpurun_out
DS1=tests/golden/data/ds1
B=oracle/_ref/gpu/ptw_batched
run() { # threads offset poll
  local s=$(date +%s%N)
  PTGPU_HOST_STATS=1 $B --raxml-info $DS1/RAxML_info.DS1 --threads $1 --polling-time $3 --lnl-offset $2 --optimization-radius 1 \
     $DS1/RAxML_bestTree.DS1 $DS1/DS1.fasta gpurun_out/good_b.tsv 2> gpurun_out/ptw_b.err
  local e=$(date +%s%N)
  echo "threads=$1 offset=$2 poll=$3 wall_ms=$(( (e - s) / 1000000 )) good=$(wc -l < gpurun_out/good_b.tsv)"; cat gpurun_out/ptw_b.err
}
run 1 -5 1
run 1 0 1
run 16 0 1
run 16 -5 1
run 16 -5 1
run 8 -5 1
