#!/bin/bash
# Round-end style check on the GPU box: GPU test suite, the headline bench line, the DS1 walk
# with the host-side wall-time accounting.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json
DS1=tests/golden/data/ds1
for i in 1 2; do
( time PTGPU_HOST_STATS=1 oracle/_ref/gpu/ptw_batched --raxml-info $DS1/RAxML_info.DS1 --threads 1 --lnl-offset -5 --optimization-radius 1 $DS1/RAxML_bestTree.DS1 $DS1/DS1.fasta gpurun_out/good_ds1.tsv ) > gpurun_out/ptw_stats_$i.log 2>&1
done
cat gpurun_out/ptw_stats_2.log
