"""pt's OptimizeBranch restated in the oracle against the compiled reference on the branches of DS5
stand-in trees (many optima at zero length), every branch in OptimizeAllBranchesOnce order; prints
(#optimisations, #results below 1e-5, #results that differ bitwise).  Needs the dump of
tests/golden/ds5_replay_local_tests.py.  (round 2: 13 328 optimisations, 1390 near zero, 0 differ)"""
import gzip,sys,os,tempfile
import numpy as np
sys.path.insert(0,'/root/repo')
from oracle.pyoracle import RefLib, Oracle
R=RefLib(); O=Oracle()
S5='/root/repo/tests/golden/data/standins/DS5'
rows=[l.rstrip('\n').split(' ',2) for l in gzip.open('/root/repo/gpurun_out/r2_ds5_dump.txt.gz','rt')]
d=tempfile.mkdtemp()
bad=[]; n=0; small=0
step=int(sys.argv[1]) if len(sys.argv)>1 else 200
for i in range(0,len(rows),step):
    nw=rows[i][2]
    f=os.path.join(d,'t.nwk'); open(f,'w').write(nw+'\n')
    p=R.open(f, S5+'/DS5.fasta', S5+'/RAxML_info.DS5')
    p.traversal_update(p.root, True)
    ev,iv,el=p.eigen(); fr,su,ra=p.model()
    recs=p.records()
    for node in p.nodes_order():
        node=int(node)
        if recs[node]['next']<0: node=recs[node]['back']
        p.traversal_update(node, False)
        st=p.sumtable(node)
        len0=p.length(node)
        want=p.optimize_branch(node)
        got,it,status=O.optimize_branch(p.sites,p.cats,st,el,ra,len0)
        n+=1
        if want<1e-5: small+=1
        if got!=want: bad.append((i,node,len0,want,got,it,status))
    p.close()
print(n, small, len(bad))
for b in bad[:30]: print(b)
