import json, sys
from collections import defaultdict
def load(p):
    agg=defaultdict(lambda:[0,0.0,0.0,0.0])
    for r in json.load(open(p)):
        key=(r['kind'],tuple(r['in']),tuple(r['out']),str(r['k']),str(r['stride']))
        a=agg[key]; a[0]+=1; a[1]+=r['ms']; a[2]+=r['gops']; a[3]+=r['mbytes']
    return agg
a=load(sys.argv[1]); b=load(sys.argv[2])
ta=sum(v[1] for v in a.values()); tb=sum(v[1] for v in b.values())
print(f"total {ta:.3f} -> {tb:.3f} ms")
for k,v in sorted(b.items(), key=lambda kv:-kv[1][1]):
    o=a.get(k)
    print(f"{k[0]:8s} {str(k[1]):16s}->{str(k[2]):16s} k{k[3]} s{k[4]} x{v[0]:<2d} {o[1] if o else 0:.3f} -> {v[1]:.3f} ms  {v[2]/v[1]:6.0f} TOPS {v[3]/v[1]:6.0f} GB/s")
