#!/bin/bash
# DS1 topography walk (BASELINE configs[1]) with the batched move tester at several wanderer counts:
# every wanderer owns a Partition = one engine + stream; their small kernels overlap on the GPU.
mkdir -p gpurun_out
DS1=tests/golden/data/ds1
B=oracle/_ref/gpu/ptw_batched
run() { # threads extra-env
  local T=$1; shift
  local s=$(date +%s%N)
  env "$@" $B --raxml-info $DS1/RAxML_info.DS1 --threads $T --polling-time ${POLL:-100} --lnl-offset -5 --optimization-radius 1 \
     --tested-trees gpurun_out/tested_$T.tsv $DS1/RAxML_bestTree.DS1 $DS1/DS1.fasta gpurun_out/good_$T.tsv 2> gpurun_out/ptw_T$T.err
  local rc=$?
  local e=$(date +%s%N)
  echo "threads=$T env=$* rc=$rc wall_ms=$(( (e - s) / 1000000 )) good=$(wc -l < gpurun_out/good_$T.tsv) tested=$(wc -l < gpurun_out/tested_$T.tsv) $(grep batched gpurun_out/ptw_T$T.err)"
}
run 1 A=1
export POLL=1
for M in 1 2 4 8; do for T in 8 16; do for rep in 1 2 3; do run $T PTGPU_COOP_SHARE_MAX=$M; done; done; done
