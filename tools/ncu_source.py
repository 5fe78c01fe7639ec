import csv, sys
from collections import Counter
rows=list(csv.reader(open(sys.argv[1])))
hi=[i for i,r in enumerate(rows) if 'Source' in r and any('Instructions Executed' in c for c in r)]
h=rows[hi[0]]
si=h.index('Source'); ie=h.index('Instructions Executed'); st=h.index('Warp Stall Sampling (All Samples)')
data=[]
for r in rows[hi[0]+1:]:
    try: data.append((int(r[st]), int(r[ie]), r[si]))
    except: pass
tots=sum(d[0] for d in data); toti=sum(d[1] for d in data)
print("total stall samples", tots, "total inst", toti)
print("--- top by stall samples")
for d in sorted(data, reverse=True)[:int(sys.argv[2]) if len(sys.argv)>2 else 22]:
    print("%5.1f%% stalls  %5.2f%% inst : %s"%(100*d[0]/tots, 100*d[1]/toti, d[2][:100]))
c=Counter()
for d in data:
    t=d[2].split()
    op=t[0] if not t[0].startswith('@') else t[1]
    c[op.split('.')[0]]+=d[1]
print("--- mix")
for k,v in c.most_common(18): print("%6.2f%% %s"%(100*v/toti,k))
