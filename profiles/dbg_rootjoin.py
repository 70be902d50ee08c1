import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, gffx_b200
from tests import _oracle
p="/tmp/rjq/plain.gff"
os.makedirs("/tmp/rjq", exist_ok=True)
gffx_b200.synth_gff(p, seed=11, n_genes=3000, quirks=False)
_oracle.build_index(p)
def op(k):
    os.environ["GFFX_WJ_KERNEL"]=k
    return gffx_b200.TreeIndexData.load_tree_index(p, device=0)
a=op("roots"); b=op("rounds")
L=int(sys.argv[1]) if len(sys.argv)>1 else 150
N=int(sys.argv[2]) if len(sys.argv)>2 else 200000
regs=a.synth_regions(5, N, L)
for mode in (2,0,1):
    ra=a.query_features(regs, mode, False); rb=b.query_features(regs, mode, False)
    print("mode",mode,"pairs",ra.n_pairs,rb.n_pairs)
    def cnt(r):
        if r.pairs is None: return {}
        u,c=np.unique(r.pairs,axis=0,return_counts=True)
        return {tuple(x):int(k) for x,k in zip(u,c)}
    ca,cb=cnt(ra),cnt(rb)
    bad=[(k,ca.get(k,0),cb.get(k,0)) for k in set(ca)|set(cb) if ca.get(k,0)!=cb.get(k,0)]
    print(" differing roots",len(bad),"of",len(cb))
    for k,x,y in sorted(bad)[:12]: print("  fid,start,end",k,"roots",x,"rounds",y)
