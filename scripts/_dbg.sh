S5=tests/golden/data/standins/DS5
rm -f /tmp/dump.txt
PTGPU_DUMP_LOCAL_EVERY=17 PTGPU_DUMP_LOCAL_TESTS=/tmp/dump.txt timeout 300 pt_b200/bin/ptw_batched --threads 1 --raxml-info $S5/RAxML_info.DS5 --lnl-offset -0.25 --optimization-radius 1 --tested-trees /tmp/t.tsv $S5/DS5.tre $S5/DS5.fasta /tmp/g.tsv 2>&1 | tail -2
wc -l /tmp/dump.txt; awk '{print $1}' /tmp/dump.txt | sort | uniq -c
gzip -c /tmp/dump.txt > gpurun_out/r2_ds5_dump.txt.gz; ls -la gpurun_out/r2_ds5_dump.txt.gz
