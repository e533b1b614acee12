#!/bin/bash
# Wall time of the DS8 stand-in walk (64 taxa x 1008 sites, 1480 visited / 115654 tested trees):
# ptw_batched with 1 / 8 / 16 wanderers on one GPU vs the reference's ptw_threads on the host cores.
mkdir -p gpurun_out
D=tests/golden/data/standins/DS8
A="--raxml-info $D/RAxML_info.DS8 --lnl-offset -1.5 --optimization-radius 0"
t() { local s=$(date +%s%N); "$@" 2> gpurun_out/walk.err; local e=$(date +%s%N); echo "wall_ms=$(( (e - s) / 1000000 )) good=$(wc -l < gpurun_out/walk_good.tsv) :: $(grep batched gpurun_out/walk.err) :: $*" | cut -c1-260; }
for T in 1 8 16; do
  t oracle/_ref/gpu/ptw_batched $A --threads $T --polling-time 1 $D/DS8.tre $D/DS8.fasta gpurun_out/walk_good.tsv
done
t oracle/_ref/gpu/ptw_batched $A --threads 16 --polling-time 1 $D/DS8.tre $D/DS8.fasta gpurun_out/walk_good.tsv
nproc
for T in 1 16; do
  t oracle/_ref/ptw_threads $A --threads $T --polling-time 1 $D/DS8.tre $D/DS8.fasta gpurun_out/walk_good.tsv
done
