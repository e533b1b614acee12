D1=tests/golden/data/ds1
for i in 1 2; do
PTGPU_HOST_STATS=1 pt_b200/bin/ptw_batched --threads 1 --polling-time 1 --raxml-info $D1/RAxML_info.DS1 --lnl-offset -5 --optimization-radius 1 --tested-trees /tmp/t.tsv $D1/RAxML_bestTree.DS1 $D1/DS1.fasta /tmp/g.tsv 2>&1 | grep "ptw_batched\]\|Partition()" 
done
