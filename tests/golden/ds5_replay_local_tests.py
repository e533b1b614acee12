"""Replays radius-1 move tests that `ptw_batched` dumped on a GPU box (PTGPU_DUMP_LOCAL_TESTS=<file>
[PTGPU_DUMP_LOCAL_EVERY=n]: "<status> <score> <moved tree, written from the move's node, 17-digit lengths>")
on the compiled reference here: FULL traversal at the move's node, OptimizeBranchNeighborhood(node, 1),
LogLikelihood -- what BranchNeighborhoodOptimizer::EvaluateMove computes -- and lists the cases whose
score differs by more than 1e-9 relative.  usage: python tests/golden/ds5_replay_local_tests.py <part> <parts>
(round 2: 33 803 sampled tests of the DS5 stand-in walk, 0 differ: profiles/r2_ds5_replay.txt)"""
import gzip,sys,os,tempfile
sys.path.insert(0,'/root/repo')
from oracle.pyoracle import RefLib
part=int(sys.argv[1]); nparts=int(sys.argv[2])
R=RefLib()
S5='/root/repo/tests/golden/data/standins/DS5'
rows=[l.rstrip('\n').split(' ',2) for l in gzip.open('/root/repo/gpurun_out/r2_ds5_dump.txt.gz','rt')]
d=tempfile.mkdtemp()
bad=[]
for i,(st,score,nw) in enumerate(rows):
    if i%nparts!=part: continue
    f=os.path.join(d,'t.nwk'); open(f,'w').write(nw+'\n')
    p=R.open(f, S5+'/DS5.fasta', S5+'/RAxML_info.DS5')
    p.traversal_update(p.root, True)
    l0=p.loglikelihood(p.root)
    p.optimize_branch_neighborhood(p.root, 1)
    p.traversal_update(p.root, False)
    l1=p.loglikelihood(p.root)
    p.close()
    g=float(score)
    if abs(g-l1)>1e-9*abs(l1): bad.append((i,g,l1,l0))
with open(f'/tmp/replay_{part}.txt','w') as f:
    for b in bad: f.write("%d %.17g %.17g %.17g\n"%b)
print(part,len(bad))
